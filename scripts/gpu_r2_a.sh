#!/bin/bash
# round 2, first GPU pass: the whole -m gpu suite, then one default bench line at N = 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/r02_a_gpu.txt
LBCHEM_VERBOSE=1 timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02_a_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r02_a_pytest.log
tail -5 gpurun_out/r02_a_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_a_bench.json 2> gpurun_out/r02_a_bench.err
echo "bench exit $?"
cat gpurun_out/r02_a_bench.json | head -c 6000
tail -5 gpurun_out/r02_a_bench.err
