#!/bin/bash
# A/B of prebuilt library variants under scripts/experiments/ab/liblbchem_<name>.so: usage gpu_ab.sh "nameA nameB" [cells] [solvers]
OUT=gpurun_out; mkdir -p $OUT
LIB=loadbalancedchemistrymodel_b200/csrc/liblbchem.so
cp $LIB /tmp/lib_orig.so
for V in $1; do
  cp scripts/experiments/ab/liblbchem_$V.so $LIB
  for S in ${3:-seulex Rosenbrock34}; do
    timeout 600 python bench.py --cells ${2:-200000} --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --solver $S > $OUT/ab_${V}_$S.json 2> $OUT/ab_${V}_$S.err
    python -c "import json;d=json.load(open('$OUT/ab_${V}_$S.json'));print('$V $S',d['value'],d['failed_cells'])" || tail -5 $OUT/ab_${V}_$S.err
  done
done
cp /tmp/lib_orig.so $LIB
