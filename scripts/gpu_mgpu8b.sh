#!/bin/bash
# usage: gpurun --gpus 8 -- 'bash scripts/gpu_mgpu8b.sh 8 TAG' : config 4 (skewed, 4 M cells, deltaTChem carried) and config 5 (DAC, 1 M cells) on N GPUs
N=${1:-8}; TAG=${2:-mg8b}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 200 $TR bench.py --gpus $N --steps 3 --warmup 3 --workload gri_skewed --cells $((4000000 / N)) --no-e2e --no-cpu-baseline > $OUT/${TAG}_bench_skewed.json 2> $OUT/${TAG}_bench_skewed.err; echo "bench skewed exit $?"
timeout 200 $TR bench.py --gpus $N --steps 3 --warmup 3 --workload gri_dac --cells $((1000000 / N)) --no-e2e --no-cpu-baseline > $OUT/${TAG}_bench_dac.json 2> $OUT/${TAG}_bench_dac.err; echo "bench dac exit $?"
for f in skewed dac; do python -c "import json;d=json.load(open('$OUT/${TAG}_bench_$f.json'));print('$f',d['n_gpus'],d['value'],d['ms_per_step'],d.get('balance_last_step'),d['gpu_launches'])" || tail -5 $OUT/${TAG}_bench_$f.err; done
