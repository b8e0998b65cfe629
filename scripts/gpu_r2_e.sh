#!/bin/bash
mkdir -p gpurun_out
run() { tag=$1; shift
  env "$@" python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced $ARGS > gpurun_out/s_$tag.json 2> gpurun_out/s_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/s_$tag.json"))
    print("$tag", round(d["value"]), "cells/s  ms", round(d["ms_per_step"],2), "failed", d["failed_cells"], "frac", round(d["roofline"]["frac"],4))
except Exception as e:
    print("$tag FAILED", e); print(open("gpurun_out/s_$tag.err").read()[-800:])
PY
}
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02_e_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r02_e_pytest.log
tail -4 gpurun_out/r02_e_pytest.log
ARGS="--workload h2_air --solver Rosenbrock34 --cells 1000000"
run h2_v3t16 X=1
run h2_v3t32 LBCHEM_V3_TPC=32
ARGS="--workload h2_air --solver seulex --cells 1000000"
run h2s_v3t16 X=1
ARGS="--workload ch4_mixing --solver Rosenbrock34 --cells 1000000"
run ch4_v3t8 X=1
run ch4_v3t16 LBCHEM_V3_TPC=16
ARGS="--workload gri_dac --solver seulex --cells 400000"
run dac_compact X=1
ARGS="--workload gri_tiled --solver seulex --cells 400000"
run nodac X=1
ARGS="--workload gri_uniform --solver seulex --cells 200000"
run gri_s X=1
ARGS="--workload gri_uniform --solver Rosenbrock34 --cells 200000"
run gri_r X=1
