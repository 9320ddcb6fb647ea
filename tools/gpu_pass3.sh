#!/bin/bash
# full GPU suite at HEAD + bench (b64) + HBM kernels + C5 shape on one GPU
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench b64 rc=$?"; tail -2 gpurun_out/bench_b64.err
python - <<'P'
import json
d=json.load(open('gpurun_out/bench_b64.json'))
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'roof',d.get('roofline',{}).get('achieved'))
for k,v in list(d['step_model']['kernel_ms_by_kind'].items())[:12]: print(f"  {k:14s} {v}")
P
timeout 600 python tools/hbm_bench.py > gpurun_out/hbm_kernels.json 2> gpurun_out/hbm_kernels.err; echo "hbm rc=$?"
python -c "
import json;h=json.load(open('gpurun_out/hbm_kernels.json'))
for k,v in h['kernels'].items(): print(k, v['us'], v['gbs'], v['frac_of_peak'])"
timeout 900 python bench.py --size 1024 --batch 64 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; echo "bench c5 rc=$?"; tail -2 gpurun_out/bench_c5_n1.err
python -c "
import json;d=json.load(open('gpurun_out/bench_c5_n1.json'));print('C5 n1',d['value'],d['ms_per_step'],d['roofline']['achieved'],d['e2e']['value'])"
