#!/bin/bash
# bench only (both solvers), no tests. usage: gpurun -- 'bash scripts/gpu_bench2.sh TAG [cells]'
TAG=${1:-q}; CELLS=${2:-200000}; OUT=gpurun_out; mkdir -p $OUT
for S in seulex Rosenbrock34; do
  timeout 600 python bench.py --cells $CELLS --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --solver $S > $OUT/${TAG}_bench_$S.json 2> $OUT/${TAG}_bench_$S.err
  python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$S.json'));print('$S',d['value'],d['roofline']['frac'],d['failed_cells'])" || tail -5 $OUT/${TAG}_bench_$S.err
done
