#!/bin/bash
# the driver's torchrun command line of bench.py at N GPUs (no tests): usage gpu_scale.sh N
mkdir -p gpurun_out
N=${1:-8}
NCCL_DEBUG=WARN timeout 1700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
echo "bench exit $?"
python - <<PY
import json
d=json.load(open("gpurun_out/r02_bench_n$N.json"))
print("value", d["value"], "e2e", d["e2e"]["value"], "ms", d["ms_per_step"])
for k in ("imbalanced","config4_gri_skewed_4M","config5_gri_dac_1M"):
    im=d.get(k)
    if im: print(k, {q: im[q] for q in ("ms_per_step_unbalanced","ms_per_step_balanced","cells_moved","nvlink_bytes_per_step","efficiency_vs_ideal","bit_identical_to_unbalanced","cells_per_s_balanced","failed_cells")}, im["e2e"]["value"])
PY
tail -2 gpurun_out/r02_bench_n$N.err
