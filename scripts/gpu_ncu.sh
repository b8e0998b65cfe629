#!/bin/bash
# one ncu --set full capture of the integration kernel. usage: gpurun -- 'bash scripts/gpu_ncu.sh TAG solver [cells]'
TAG=${1:-n}; S=${2:-seulex}; CELLS=${3:-20000}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 python bench.py --cells $CELLS --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --solver $S > $OUT/${TAG}_plain.json 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:integrate -s 1 -c 1 -f -o $OUT/${TAG}_full \
  python bench.py --cells $CELLS --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --solver $S > $OUT/${TAG}_ncu.log 2>&1; echo "ncu exit $?"
