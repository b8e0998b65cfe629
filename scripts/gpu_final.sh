#!/bin/bash
# last run of the round: HEAD check (tests, smoke, both arms at defaults), Rosenbrock34 and config 5 at 1 M cells,
# one ncu --set full capture of the Rosenbrock34 instantiation at 100 k cells
TAG=${1:-r01_final2}; OUT=gpurun_out; mkdir -p $OUT
bash scripts/gpu_head_check.sh $TAG
timeout 400 python bench.py --solver Rosenbrock34 > $OUT/${TAG}_bench_ros.json 2> $OUT/${TAG}_bench_ros.err; echo "bench ros exit $?"
timeout 300 python bench.py --workload gri_dac > $OUT/${TAG}_bench_dac.json 2> $OUT/${TAG}_bench_dac.err; echo "bench dac exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:integrate -s 1 -c 1 -f -o $OUT/${TAG}_ros_full \
  python bench.py --cells 100000 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --solver Rosenbrock34 > $OUT/${TAG}_ncu_ros_full.log 2>&1; echo "ncu full exit $?"
