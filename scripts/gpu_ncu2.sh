#!/bin/bash
# ncu --set full captures (source-level): usage gpu_ncu2.sh TAG "bench args" [TAG2 "bench args2"]
OUT=gpurun_out; mkdir -p $OUT
while [ $# -ge 2 ]; do
  TAG=$1; A=$2; shift 2
  CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced $A"
  timeout 300 $CMD > $OUT/${TAG}_plain.json 2> $OUT/${TAG}_plain.err &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:integrate -s 3 -c 1 -f -o $OUT/${TAG}_full $CMD > $OUT/${TAG}_ncu.log 2>&1
  echo "$TAG ncu exit $?"
done
