#!/bin/bash
# last check of the round: whole -m gpu suite, smoke, pivoting 0 / 1 pair at 200 k cells (same box)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r02_pytest_gpu.log
tail -2 gpurun_out/r02_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/r02_smoke.log
for s in seulex Rosenbrock34; do for piv in 0 1; do
  python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced --cells 200000 --solver $s --pivoting $piv > gpurun_out/last_${s}_p$piv.json 2> gpurun_out/last_${s}_p$piv.err
  python -c "import json; d=json.load(open('gpurun_out/last_${s}_p$piv.json')); print('$s pivoting $piv', round(d['value']), 'cells/s')"
done; done
