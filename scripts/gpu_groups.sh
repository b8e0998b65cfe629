#!/bin/bash
for G in 1 2 3 4; do
  LBCHEM_GROUPS=$G timeout 600 python bench.py --cells 100000 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --solver seulex > gpurun_out/g$G.json 2>/dev/null
  python -c "import json;d=json.load(open('gpurun_out/g$G.json'));print('groups $G',d['value'])"
done
