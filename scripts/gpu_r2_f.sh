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
timeout 900 python -m pytest tests -m gpu -x -q -s -k "dac or full_size or two_ranks" > gpurun_out/r02_f_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r02_f_pytest.log
tail -4 gpurun_out/r02_f_pytest.log
ARGS="--workload gri_dac --solver seulex --cells 400000"
run dac2 X=1
ARGS="--workload gri_tiled --solver seulex --cells 400000"
run nodac2 X=1
ARGS="--workload gri_dac --solver Rosenbrock34 --cells 400000"
run dac2r X=1
ARGS="--workload gri_tiled --solver Rosenbrock34 --cells 400000"
run nodac2r X=1
