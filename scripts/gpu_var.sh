#!/bin/bash
# A/B of LBCHEM_VARIANT values in one call (same box): usage gpu_var.sh "0 1" ; parity subset runs with the LAST variant
mkdir -p gpurun_out
VARS=${1:-"0 1"}
LAST=${VARS##* }
LBCHEM_VARIANT=$LAST timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s -k "pivoting or fixture_parity or h2_air or ch4_one or dac_parity or probes or tc_kernel or edge" > gpurun_out/var_pytest.log 2>&1
echo "pytest exit $? (variant $LAST)" >> gpurun_out/var_pytest.log
tail -3 gpurun_out/var_pytest.log
run() { tag=$1; shift
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced "$@" > gpurun_out/var_$tag.json 2> gpurun_out/var_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/var_$tag.json"))
    print("$tag", round(d["value"]), "cells/s  ms", round(d["ms_per_step"],2), "failed", d["failed_cells"], "frac", round(d["roofline"]["frac"],4))
except Exception as e:
    print("$tag FAILED", e); print(open("gpurun_out/var_$tag.err").read()[-800:])
PY
}
for rep in 1 2; do
for v in $VARS; do
  export LBCHEM_VARIANT=$v
  run s_v${v}_$rep --workload gri_uniform --solver seulex --cells 200000
  run r_v${v}_$rep --workload gri_uniform --solver Rosenbrock34 --cells 200000
done; done
for v in $VARS; do
  export LBCHEM_VARIANT=$v
  run h2_v$v --workload h2_air --solver Rosenbrock34 --cells 1000000
  run dac_v$v --workload gri_dac --solver seulex --cells 400000
done
