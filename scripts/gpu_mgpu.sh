#!/bin/bash
# usage: gpurun --gpus N -- 'bash scripts/gpu_mgpu.sh N TAG'
N=${1:-2}; TAG=${2:-mg}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR scripts/mgpu_check.py 20000 > $OUT/${TAG}_check.log 2>&1; echo "check exit $?"; grep -E "MGPU_CHECK|identical=False|rror" $OUT/${TAG}_check.log | head; grep "rank 0 step" $OUT/${TAG}_check.log
timeout 900 $TR bench.py --gpus $N --steps 3 --warmup 3 --cells 400000 --workload gri_skewed --no-e2e --no-cpu-baseline > $OUT/${TAG}_bench_skewed.json 2> $OUT/${TAG}_bench_skewed.err; echo "bench skewed exit $?"
timeout 900 $TR bench.py --gpus $N --steps 3 --warmup 3 --cells 400000 --no-e2e --no-cpu-baseline > $OUT/${TAG}_bench_uniform.json 2> $OUT/${TAG}_bench_uniform.err; echo "bench uniform exit $?"
for f in skewed uniform; do python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$f.json'));print('$f',d['n_gpus'],d['value'],d['ms_per_step'],d.get('balance_last_step'))" || tail -5 $OUT/${TAG}_bench_$f.err; done
