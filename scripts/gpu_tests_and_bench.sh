#!/bin/bash
# round 2, first GPU pass: the whole -m gpu suite, then one default bench line at N = 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/tb_gpu.txt
LBCHEM_VERBOSE=1 timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/tb_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/tb_pytest.log
tail -5 gpurun_out/tb_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/tb_bench.json 2> gpurun_out/tb_bench.err
echo "bench exit $?"
cat gpurun_out/tb_bench.json | head -c 6000
tail -5 gpurun_out/tb_bench.err
