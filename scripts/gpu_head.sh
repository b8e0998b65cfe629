#!/bin/bash
# bench lines of the final commit: default line (config 3), config 5 with and without DAC
O=gpurun_out; mkdir -p $O
timeout 900 python bench.py --steps 10 --warmup 3 > $O/r02_bench_n1_seulex.json 2> $O/r02_bench_n1_seulex.err; echo "bench exit $?"
timeout 600 python bench.py --steps 5 --warmup 3 --workload gri_dac --no-imbalanced > $O/r02_bench_n1_dac.json 2> $O/r02_bench_n1_dac.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload gri_tiled --no-imbalanced > $O/r02_bench_n1_tiled.json 2> $O/r02_bench_n1_tiled.err
for f in seulex dac tiled; do python - <<PY
import json
d=json.load(open("$O/r02_bench_n1_$f.json")); print("$f", round(d["value"]), "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"],4), "cpu", round(d["cpu_baseline"]["value"]), "traffic", d["roofline"]["traffic"])
PY
done
