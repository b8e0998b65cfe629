#!/bin/bash
# config 5 (GRI + DAC) against the same cells with the full mechanism, one GPU
mkdir -p gpurun_out
N=${1:-1000000}
for w in gri_dac gri_tiled; do
  LBCHEM_VERBOSE=1 timeout 900 python bench.py --workload $w --cells $N --steps 3 --warmup 3 > gpurun_out/dacb_$w.json 2> gpurun_out/dacb_$w.err
  echo "$w exit $?"
  python - <<PY
import json
d=json.load(open("gpurun_out/dacb_$w.json"))
print("$w", d["value"], d["e2e"]["value"], d["cpu_baseline"], d["stats_last_step"])
PY
done
