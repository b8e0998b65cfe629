#!/bin/bash
# small mechanisms: which kernel / group width, then the loopback + remaining tests
mkdir -p gpurun_out
run() { # tag env... -- args
  tag=$1; shift
  env "$@" python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced $ARGS > gpurun_out/s_$tag.json 2> gpurun_out/s_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/s_$tag.json"))
    print("$tag", round(d["value"]), "cells/s  ms", round(d["ms_per_step"],2), "failed", d["failed_cells"], "frac", round(d["roofline"]["frac"],4))
except Exception as e:
    print("$tag FAILED", e); print(open("gpurun_out/s_$tag.err").read()[-600:])
PY
}
ARGS="--workload h2_air --solver Rosenbrock34 --cells 1000000"
run h2_v3 LBCHEM_SMALL_KERNEL=v3
run h2_g8 LBCHEM_TPC=8
run h2_g16 LBCHEM_TPC=16
run h2_g32 LBCHEM_TPC=32
ARGS="--workload h2_air --solver seulex --cells 1000000"
run h2s_v3 LBCHEM_SMALL_KERNEL=v3
run h2s_g16 LBCHEM_TPC=16
ARGS="--workload ch4_mixing --solver Rosenbrock34 --cells 1000000"
run ch4_v3 LBCHEM_SMALL_KERNEL=v3
run ch4_g8 LBCHEM_TPC=8
run ch4_g16 LBCHEM_TPC=16
ARGS="--workload ch4_mixing --solver EulerImplicit --cells 1000000"
run ch4e_g8 LBCHEM_TPC=8
run ch4e_g16 LBCHEM_TPC=16
run ch4e_g32 LBCHEM_TPC=32
timeout 1500 python -m pytest tests -m gpu -x -q -s -k "two_ranks or h2_air or ch4 or edge" > gpurun_out/r02_d_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r02_d_pytest.log
tail -4 gpurun_out/r02_d_pytest.log
grep -n "step [0-9]:" gpurun_out/r02_d_pytest.log | cut -c1-300
