#!/bin/bash
mkdir -p gpurun_out

timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s -k "pivoting or fixture_parity or h2_air or ch4_one or dac_parity or probes" > gpurun_out/piv_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/piv_pytest.log
tail -3 gpurun_out/piv_pytest.log
run() { tag=$1; shift
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced "$@" > gpurun_out/piv_$tag.json 2> gpurun_out/piv_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/piv_$tag.json"))
    print("$tag", round(d["value"]), "cells/s  ms", round(d["ms_per_step"],2), "failed", d["failed_cells"], "frac", round(d["roofline"]["frac"],4))
except Exception as e:
    print("$tag FAILED", e); print(open("gpurun_out/piv_$tag.err").read()[-800:])
PY
}
run s_p1 --workload gri_uniform --solver seulex --cells 200000 --pivoting 1
run r_p1 --workload gri_uniform --solver Rosenbrock34 --cells 200000 --pivoting 1
run h2_p1 --workload h2_air --solver Rosenbrock34 --cells 1000000 --pivoting 1
run ch4_p1 --workload ch4_mixing --solver Rosenbrock34 --cells 1000000 --pivoting 1
LBCHEM_VERBOSE=1 run dac_p1 --workload gri_dac --solver seulex --cells 400000 --pivoting 1; grep "v3 kernel:" gpurun_out/piv_dac_p1.err | tail -1
run s_p0 --workload gri_uniform --solver seulex --cells 200000 --pivoting 0
