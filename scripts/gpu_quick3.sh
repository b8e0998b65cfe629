#!/bin/bash
# tight-timeout iteration: v3 + DAC parity, then short benches. A wrong kernel must not be able to burn the GPU budget.
OUT=gpurun_out; TAG=${1:-q3}; mkdir -p $OUT
timeout 150 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "v3 or dac" > $OUT/${TAG}_pytest.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -3 $OUT/${TAG}_pytest.log
[ $rc -ne 0 ] && exit 1
for S in seulex Rosenbrock34; do
  timeout 100 python bench.py --cells 200000 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --solver $S > $OUT/${TAG}_bench_$S.json 2> $OUT/${TAG}_bench_$S.err
  python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$S.json'));print('$S',d['value'],d['failed_cells'])" || tail -5 $OUT/${TAG}_bench_$S.err
done
for W in gri_dac gri_tiled; do
  timeout 100 python bench.py --cells 400000 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --workload $W > $OUT/${TAG}_bench_$W.json 2> $OUT/${TAG}_bench_$W.err
  python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$W.json'));print('$W',d['value'],d['failed_cells'])" || tail -5 $OUT/${TAG}_bench_$W.err
done
