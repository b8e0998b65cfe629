#!/bin/bash
# phase timing of the v3 kernel: rebuild with -DV3_TIMING on the box, run a small bench per solver, restore is not needed (box is scratch)
cd loadbalancedchemistrymodel_b200/csrc && rm -f lbchem_kernels_v3.o && make NVCC="nvcc -DV3_TIMING" -j4 > /dev/null 2>&1; cd ../..
for S in seulex Rosenbrock34; do
  LBCHEM_TIMING=1 timeout 600 python bench.py --cells ${1:-50000} --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --solver $S 2>&1 | grep "v3 timing" | tail -1
done
