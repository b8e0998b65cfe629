#!/bin/bash
# confirmation of the round's HEAD on a fresh box: GPU tests, smoke, the driver's two default bench lines
TAG=${1:-r01_head}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 python -m pytest tests -m gpu -x -q -s > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $OUT/${TAG}_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $OUT/${TAG}_smoke.log
timeout 300 python bench.py --impl reference > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "reference exit $?"
timeout 300 python bench.py > $OUT/${TAG}_bench_seulex.json 2> $OUT/${TAG}_bench_seulex.err; echo "bench seulex exit $?"
cut -c1-300 $OUT/${TAG}_bench_seulex.json
