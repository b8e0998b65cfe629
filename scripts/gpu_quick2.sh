#!/bin/bash
# parity tests + quick benches (uniform seulex / Rosenbrock34, config-5 input with and without DAC)
OUT=gpurun_out; TAG=${1:-q2}; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest.log
for S in seulex Rosenbrock34; do
  timeout 600 python bench.py --cells 200000 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --solver $S > $OUT/${TAG}_bench_$S.json 2> $OUT/${TAG}_bench_$S.err
  python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$S.json'));print('$S',d['value'],d['failed_cells'])" || tail -5 $OUT/${TAG}_bench_$S.err
done
for W in gri_dac gri_tiled; do
  timeout 600 python bench.py --cells 400000 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --workload $W > $OUT/${TAG}_bench_$W.json 2> $OUT/${TAG}_bench_$W.err
  python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$W.json'));print('$W',d['value'],d['failed_cells'])" || tail -5 $OUT/${TAG}_bench_$W.err
done
