#!/bin/bash
mkdir -p gpurun_out
run() { tag=$1; shift
  env LBCHEM_VERBOSE=1 "$@" python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced $ARGS > gpurun_out/s_$tag.json 2> gpurun_out/s_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/s_$tag.json"))
    print("$tag", round(d["value"]), "cells/s  ms", round(d["ms_per_step"],2), "failed", d["failed_cells"], "frac", round(d["roofline"]["frac"],4))
except Exception as e:
    print("$tag FAILED", e); print(open("gpurun_out/s_$tag.err").read()[-800:])
PY
}
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/ab_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/ab_pytest.log
tail -4 gpurun_out/ab_pytest.log
ARGS="--workload gri_uniform --solver seulex --cells 200000"
run i_s X=1
ARGS="--workload gri_uniform --solver Rosenbrock34 --cells 200000"
run i_r X=1
ARGS="--workload gri_dac --solver seulex --cells 400000"
run i_dac X=1
ARGS="--workload gri_tiled --solver seulex --cells 400000"
run i_nodac X=1
ARGS="--workload h2_air --solver Rosenbrock34 --cells 1000000"
run i_h2 X=1
ARGS="--workload ch4_mixing --solver Rosenbrock34 --cells 1000000"
run i_ch4 X=1
# where DAC pays: hot cells only (config 3 input) and cold cells only, with and without the reduction
ARGS="--workload gri_uniform --solver seulex --cells 200000 --dac"
run hot_dac X=1
ARGS="--workload gri_uniform --solver seulex --cells 200000"
run hot_nodac LBCHEM_GROUPS=6
