#!/bin/bash
# remaining new GPU tests (1 GPU of the box), then the driver's N = 2 command line with a short step count
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -s -k "tight or pipeline or cost_ordered or cells_to_send or two_ranks or driver_on_one" > gpurun_out/r02_c_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/r02_c_pytest.log
tail -4 gpurun_out/r02_c_pytest.log
grep -n "step [0-9]:" gpurun_out/r02_c_pytest.log | cut -c1-300
NCCL_DEBUG=WARN timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_c_bench2.json 2> gpurun_out/r02_c_bench2.err
echo "bench2 exit $?"
tail -c 5000 gpurun_out/r02_c_bench2.json
tail -15 gpurun_out/r02_c_bench2.err
