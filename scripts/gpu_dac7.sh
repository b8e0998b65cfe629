#!/bin/bash
# DAC with the enabled-reaction list in the matrix tail (seven cells per SM) against the separate block (six): tests + same-box pair
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s -k "dac or pivoting or full_size or two_ranks or miss_list" > gpurun_out/dac7_pytest.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/dac7_pytest.log; tail -2 gpurun_out/dac7_pytest.log
run() { tag=$1; shift
  LBCHEM_VERBOSE=1 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-imbalanced "$@" > gpurun_out/dac7_$tag.json 2> gpurun_out/dac7_$tag.err
  python -c "import json; d=json.load(open('gpurun_out/dac7_$tag.json')); print('$tag', round(d['value']), 'cells/s failed', d['failed_cells'])"
  grep "v3 kernel:" gpurun_out/dac7_$tag.err | tail -1
}
for rep in 1 2; do
LBCHEM_RL_SEPARATE=1 run sep$rep --workload gri_dac --solver seulex --cells 400000
run tail$rep --workload gri_dac --solver seulex --cells 400000
done
run full --workload gri_tiled --solver seulex --cells 400000
LBCHEM_RL_SEPARATE=1 run hot_sep --workload gri_uniform --solver seulex --cells 200000 --dac
run hot_tail --workload gri_uniform --solver seulex --cells 200000 --dac
