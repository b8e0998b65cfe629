#!/bin/bash
# parity subset on the in-tree build, then same-box A/B of builds (scripts/gpu_libs.sh)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s -k "pivoting or fixture_parity or h2_air or ch4_one or dac_parity or probes or tc_kernel or edge" > gpurun_out/var_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/var_pytest.log
tail -3 gpurun_out/var_pytest.log
bash scripts/gpu_libs.sh "$1"
