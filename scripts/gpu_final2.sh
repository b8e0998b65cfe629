#!/bin/bash
# N = 2: the whole -m gpu suite on one GPU, then the driver's torchrun command line of both arms
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r02_pytest_gpu.log
tail -2 gpurun_out/r02_pytest_gpu.log
NCCL_DEBUG=WARN timeout 1700 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_bench_n2.json 2> gpurun_out/r02_bench_n2.err
echo "bench exit $?"
python - <<PY
import json
d=json.load(open("gpurun_out/r02_bench_n2.json"))
print("value", d["value"], "e2e", d["e2e"]["value"], "ms", d["ms_per_step"])
im=d.get("imbalanced")
if im: print({q: im[q] for q in ("ms_per_step_unbalanced","ms_per_step_balanced","cells_moved","nvlink_bytes_per_step","efficiency_vs_ideal","bit_identical_to_unbalanced","cells_per_s_balanced","failed_cells")}, im["e2e"]["value"])
PY
tail -3 gpurun_out/r02_bench_n2.err
