#!/bin/bash
# same-box A/B of several builds of liblbchem.so (csrc/variants/liblbchem_<TAG>.so): usage gpu_libs.sh "A B C" [variant env per tag e.g. B:1]
mkdir -p gpurun_out
run() { tag=$1; shift
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced "$@" > gpurun_out/lib_$tag.json 2> gpurun_out/lib_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/lib_$tag.json"))
    print("$tag", round(d["value"]), "cells/s  ms", round(d["ms_per_step"],2), "failed", d["failed_cells"])
except Exception as e:
    print("$tag FAILED", e); print(open("gpurun_out/lib_$tag.err").read()[-600:])
PY
}
for rep in 1 2; do
for t in $1; do
  tag=${t%%:*}; var=0; [[ "$t" == *:* ]] && var=${t##*:}
  export LBCHEM_LIB=$PWD/loadbalancedchemistrymodel_b200/csrc/variants/liblbchem_$tag.so LBCHEM_VARIANT=$var
  run ${tag}v${var}_s$rep --workload gri_uniform --solver seulex --cells 200000
done; done
for t in $1; do
  tag=${t%%:*}; var=0; [[ "$t" == *:* ]] && var=${t##*:}
  export LBCHEM_LIB=$PWD/loadbalancedchemistrymodel_b200/csrc/variants/liblbchem_$tag.so LBCHEM_VARIANT=$var
  run ${tag}v${var}_r --workload gri_uniform --solver Rosenbrock34 --cells 200000
  run ${tag}v${var}_h2 --workload h2_air --solver Rosenbrock34 --cells 1000000
done
