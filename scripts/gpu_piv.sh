#!/bin/bash
# diagonal-pivot path: GPU tests, then A/B of lbchem_options.pivoting on the headline workloads (200 k cells)
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q -s > gpurun_out/piv_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/piv_pytest.log
tail -5 gpurun_out/piv_pytest.log
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
for piv in 0 1; do
  run s_p$piv --workload gri_uniform --solver seulex --cells 200000 --pivoting $piv
  run r_p$piv --workload gri_uniform --solver Rosenbrock34 --cells 200000 --pivoting $piv
  run dac_p$piv --workload gri_dac --solver seulex --cells 400000 --pivoting $piv
  run h2_p$piv --workload h2_air --solver Rosenbrock34 --cells 1000000 --pivoting $piv
  run ch4_p$piv --workload ch4_mixing --solver Rosenbrock34 --cells 1000000 --pivoting $piv
done
run tiled_p1 --workload gri_tiled --solver seulex --cells 400000 --pivoting 1
