#!/bin/bash
# A/B of lbchem_options.pivoting (200 k cells), then one full ncu capture of the seulex kernel with the diagonal path
mkdir -p gpurun_out
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
  run h2_p$piv --workload h2_air --solver Rosenbrock34 --cells 1000000 --pivoting $piv
  run ch4_p$piv --workload ch4_mixing --solver Rosenbrock34 --cells 1000000 --pivoting $piv
done
bash scripts/gpu_ncu2.sh piv1 "--workload gri_uniform --solver seulex --cells 100000 --pivoting 1"
