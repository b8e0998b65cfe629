#!/bin/bash
# DAC overhead vs savings: config-5 input with tolerance 1e-300 (nothing removed: pure overhead), 1e-4, 1e-2, and no DAC
OUT=gpurun_out; mkdir -p $OUT
for TOL in 1e-300 1e-4 1e-2; do
  LBCHEM_DAC_TOL=$TOL timeout 100 python bench.py --cells 400000 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --workload gri_dac > $OUT/dtol_$TOL.json 2> $OUT/dtol_$TOL.err
  python -c "import json;d=json.load(open('$OUT/dtol_$TOL.json'));print('tol $TOL',d['value'],d['stats_last_step']['n_attempts'])" || tail -3 $OUT/dtol_$TOL.err
done
timeout 100 python bench.py --cells 400000 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --workload gri_tiled > $OUT/dtol_none.json 2> $OUT/dtol_none.err
python -c "import json;d=json.load(open('$OUT/dtol_none.json'));print('no dac',d['value'],d['stats_last_step']['n_attempts'])"
