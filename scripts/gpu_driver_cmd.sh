#!/bin/bash
# the driver's own N>1 command lines (defaults: e2e + cpu_baseline on), both arms
N=${1:-2}; OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 250 $TR bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $OUT/drv_ref_$N.json 2> $OUT/drv_ref_$N.err; echo "reference exit $?"; wc -l $OUT/drv_ref_$N.json
timeout 300 $TR bench.py --gpus $N --steps 3 --warmup 3 > $OUT/drv_b200_$N.json 2> $OUT/drv_b200_$N.err; echo "b200 exit $?"; wc -l $OUT/drv_b200_$N.json
python -c "import json;d=json.load(open('$OUT/drv_b200_$N.json'));print(d['n_gpus'],d['value'],d['e2e']['value'],d['cpu_baseline']['value'],d['gpu_launches'],d['clocks'])"
python -c "import json;d=json.load(open('$OUT/drv_ref_$N.json'));print(d['impl'],d['value'])"
