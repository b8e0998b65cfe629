#!/bin/bash
# full-size evidence for profiles/: tests, both bench lines at 1M cells, reference arm, launch list at 1M, ncu --set full at 100k cells
TAG=${1:-r01_v3final}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 python -m pytest tests -m gpu -x -q -s > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $OUT/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $OUT/${TAG}_smoke.log
timeout 300 python bench.py > $OUT/${TAG}_bench_seulex.json 2> $OUT/${TAG}_bench_seulex.err; echo "bench seulex exit $?"
timeout 400 python bench.py --solver Rosenbrock34 > $OUT/${TAG}_bench_ros.json 2> $OUT/${TAG}_bench_ros.err; echo "bench ros exit $?"
timeout 300 python bench.py --workload gri_dac > $OUT/${TAG}_bench_dac.json 2> $OUT/${TAG}_bench_dac.err; echo "bench dac exit $?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "reference exit $?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:integrate -s 1 -c 1 -f -o $OUT/${TAG}_full \
  python bench.py --cells 100000 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
cut -c1-300 $OUT/${TAG}_bench_seulex.json
