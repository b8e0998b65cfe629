#!/bin/bash
# quick GPU iteration: parity tests + one small bench per solver. usage: gpurun -- 'bash scripts/gpu_quick.sh TAG [cells]'
TAG=${1:-q}; CELLS=${2:-200000}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest.log
for S in seulex Rosenbrock34; do
  LBCHEM_VERBOSE=1 timeout 600 python bench.py --cells $CELLS --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --solver $S > $OUT/${TAG}_bench_$S.json 2> $OUT/${TAG}_bench_$S.err; echo "bench $S exit $?"
  grep -m1 "lbchem\] v3" $OUT/${TAG}_bench_$S.err
  python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$S.json'));print('$S',d['value'],d['roofline']['frac'],d['failed_cells'])" || tail -5 $OUT/${TAG}_bench_$S.err
done
