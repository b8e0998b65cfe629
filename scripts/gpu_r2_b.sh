#!/bin/bash
mkdir -p gpurun_out
bash scripts/gpu_quick.sh dterm
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02_b_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r02_b_pytest.log
tail -4 gpurun_out/r02_b_pytest.log
