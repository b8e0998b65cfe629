#!/bin/bash
# DAC-reduced integration: parity test + timing (reduced vs full) on one GPU.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s -k "dac" > gpurun_out/dac_pytest.log 2>&1
echo "pytest exit $?"; tail -15 gpurun_out/dac_pytest.log
