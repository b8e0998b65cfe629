#!/bin/bash
# quick A/B: 200 k cells, both solvers, device-resident value only; tag = $1
mkdir -p gpurun_out
for s in seulex Rosenbrock34; do
  python bench.py --cells 200000 --steps 3 --warmup 3 --solver $s --no-e2e --no-cpu-baseline --no-imbalanced > gpurun_out/q_${1}_$s.json 2> gpurun_out/q_${1}_$s.err
  python - <<PY
import json
d=json.load(open("gpurun_out/q_${1}_$s.json"))
print("${1}", "$s", round(d["value"]), "cells/s  frac", round(d["roofline"]["frac"],4), "failed", d["failed_cells"])
PY
done
