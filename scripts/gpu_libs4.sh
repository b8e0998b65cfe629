#!/bin/bash
# same-box A/B of builds on five workloads: usage gpu_libs4.sh "P0 N1"
mkdir -p gpurun_out
run() { tag=$1; shift
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced "$@" > gpurun_out/lib_$tag.json 2> gpurun_out/lib_$tag.err
  python -c "import json; d=json.load(open('gpurun_out/lib_$tag.json')); print('$tag', round(d['value']), 'cells/s failed', d['failed_cells'])" || tail -5 gpurun_out/lib_$tag.err
}
for t in $1; do
  export LBCHEM_LIB=$PWD/loadbalancedchemistrymodel_b200/csrc/variants/liblbchem_$t.so
  run ${t}_s --workload gri_uniform --solver seulex --cells 200000
  run ${t}_r --workload gri_uniform --solver Rosenbrock34 --cells 200000
  run ${t}_tiled --workload gri_tiled --solver seulex --cells 400000
  run ${t}_dac --workload gri_dac --solver seulex --cells 400000
  run ${t}_h2 --workload h2_air --solver Rosenbrock34 --cells 1000000
done
