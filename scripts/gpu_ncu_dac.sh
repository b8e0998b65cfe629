#!/bin/bash
# one full ncu capture of the DAC instantiation of the v3 kernel (config-5 input, 60 k cells)
OUT=gpurun_out; mkdir -p $OUT
timeout 500 ncu --set full --clock-control none --import-source on -k regex:integrate -s 1 -c 1 -f -o $OUT/r01_dac_full \
  python bench.py --workload gri_dac --cells 60000 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/r01_dac_ncu.log 2>&1; echo "ncu exit $?"
