#!/bin/bash
# the no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last), then the
# two small-mechanism bench lines after the 512-thread change
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/noalloc_alone.log 2>&1; echo "alone exit $?" | tee -a $O/noalloc_alone.log
tail -5 $O/noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/r02_pytest_gpu.log
tail -4 $O/r02_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --workload h2_air --solver Rosenbrock34 --no-imbalanced > $O/r02_bench_n1_h2.json 2> $O/r02_bench_n1_h2.err
timeout 600 python bench.py --steps 5 --warmup 3 --workload ch4_mixing --solver Rosenbrock34 --no-imbalanced > $O/r02_bench_n1_ch4_ros.json 2> $O/r02_bench_n1_ch4_ros.err
for f in $O/r02_bench_n1_h2.json $O/r02_bench_n1_ch4_ros.json; do python - <<PY
import json
try:
    d=json.load(open("$f")); print("$f".split("/")[-1], round(d["value"]), "e2e", d.get("e2e") and round(d["e2e"]["value"]), "frac", d.get("roofline") and round(d["roofline"]["frac"],4), "cpu", d.get("cpu_baseline") and round(d["cpu_baseline"]["value"]))
except Exception as e: print("$f", "FAILED", e)
PY
done
