#!/bin/bash
# One gpurun call: GPU parity tests, a bench line per solver, the ncu launch list and one full capture.
# usage: gpurun --timeout 2400 -- 'bash scripts/gpu_check.sh TAG'
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > $OUT/${TAG}_gpu.txt 2>&1
nproc >> $OUT/${TAG}_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q -s > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
timeout 900 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_seulex.json 2> $OUT/${TAG}_bench_seulex.err; echo "bench seulex exit $?"
timeout 900 python bench.py --steps 3 --warmup 3 --solver Rosenbrock34 > $OUT/${TAG}_bench_ros.json 2> $OUT/${TAG}_bench_ros.err; echo "bench ros exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --cells 100000 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:integrate -s 1 -c 1 -f -o $OUT/${TAG}_full \
  python bench.py --cells 20000 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
tail -3 $OUT/${TAG}_pytest.log
cat $OUT/${TAG}_bench_seulex.json $OUT/${TAG}_bench_ros.json
