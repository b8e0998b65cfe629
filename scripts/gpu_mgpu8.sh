#!/bin/bash
# usage: gpurun --gpus N -- 'bash scripts/gpu_mgpu8.sh N TAG'
N=${1:-8}; TAG=${2:-mg8}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 300 $TR scripts/mgpu_check.py 20000 > $OUT/${TAG}_check.log 2>&1; echo "check exit $?"; grep -E "MGPU_CHECK" $OUT/${TAG}_check.log; grep "step 2" $OUT/${TAG}_check.log | cut -c1-200
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_bench_uniform.json 2> $OUT/${TAG}_bench_uniform.err; echo "bench uniform exit $?"
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 --workload gri_skewed --no-e2e --no-cpu-baseline > $OUT/${TAG}_bench_skewed.json 2> $OUT/${TAG}_bench_skewed.err; echo "bench skewed exit $?"
for f in uniform skewed; do python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$f.json'));print('$f',d['n_gpus'],d['value'],d['ms_per_step'],d.get('balance_last_step'),d['gpu_launches'])" || tail -5 $OUT/${TAG}_bench_$f.err; done
