#!/bin/bash
# The evidence run behind profiles/r02_*: GPU tests, smoke, the default bench line and the reference arm, the other
# BASELINE configs at 1 M cells, then ncu: launch list of the default command, DRAM traffic at 1 M cells, one full capture.
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv,noheader > $O/r02_gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -s > $O/r02_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/r02_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/r02_smoke.log 2>&1; echo "smoke exit $?" | tee -a $O/r02_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/r02_bench_n1_seulex.json 2> $O/r02_bench_n1_seulex.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_bench_n1_reference.json 2> $O/r02_bench_n1_reference.err
timeout 900 python bench.py --steps 5 --warmup 3 --solver Rosenbrock34 --no-imbalanced > $O/r02_bench_n1_rosenbrock34.json 2> $O/r02_bench_n1_rosenbrock34.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload gri_dac --no-imbalanced > $O/r02_bench_n1_dac.json 2> $O/r02_bench_n1_dac.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload gri_tiled --no-imbalanced > $O/r02_bench_n1_tiled.json 2> $O/r02_bench_n1_tiled.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload h2_air --solver Rosenbrock34 --no-imbalanced > $O/r02_bench_n1_h2.json 2> $O/r02_bench_n1_h2.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload ch4_mixing --solver EulerImplicit --no-imbalanced > $O/r02_bench_n1_ch4_euler.json 2> $O/r02_bench_n1_ch4_euler.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload ch4_mixing --solver Rosenbrock34 --no-imbalanced > $O/r02_bench_n1_ch4_ros.json 2> $O/r02_bench_n1_ch4_ros.err
for f in $O/r02_bench_n1_*.json; do python - <<PY
import json
try:
    d=json.load(open("$f")); print("$f".split("/")[-1], round(d["value"]), "e2e", d.get("e2e") and round(d["e2e"]["value"]), "frac", d.get("roofline") and round(d["roofline"]["frac"],4), "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"]))
except Exception as e: print("$f", "FAILED", e)
PY
done
# ---- ncu (one tool per call): launch list of the default command, DRAM traffic at 1 M cells, full capture at 100 k cells
SHORT="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced"
$SHORT > $O/r02_plain1.json 2>/dev/null &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv --log-file $O/r02_ncu_launches_seulex.csv $SHORT > $O/r02_ncu1.log 2>&1
$SHORT --solver Rosenbrock34 > $O/r02_plain2.json 2>/dev/null &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv --log-file $O/r02_ncu_launches_rosenbrock34.csv $SHORT --solver Rosenbrock34 > $O/r02_ncu2.log 2>&1
F="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced --cells 100000"
$F > $O/r02_plain3.json 2>/dev/null &&
ncu --set full --clock-control none --import-source on -k regex:integrate -s 3 -c 1 -f -o $O/r02_final_seulex_full $F > $O/r02_ncu3.log 2>&1
echo "ncu exit $?"
