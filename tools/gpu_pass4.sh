#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for SIDE in 0 1; do
  B2SEG_SIDE=$SIDE timeout 900 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64_side$SIDE.json 2> gpurun_out/bench_b64_side$SIDE.err; echo "bench side=$SIDE rc=$?"; tail -2 gpurun_out/bench_b64_side$SIDE.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_b64_side$SIDE.json'));print('side$SIDE value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'loss',d['final_loss'])"
done
timeout 600 python tools/hbm_bench.py > gpurun_out/hbm_kernels.json 2> gpurun_out/hbm_kernels.err; echo "hbm rc=$?"
python -c "
import json;h=json.load(open('gpurun_out/hbm_kernels.json'))
for k,v in h['kernels'].items():
    if 'seg' in k: print(k, v['us'], v['gbs'], v['frac_of_peak'])"
