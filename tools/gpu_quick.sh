#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench b64 rc=$?"; tail -2 gpurun_out/bench_b64.err
python - <<'P'
import json
d=json.load(open('gpurun_out/bench_b64.json'))
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'roof',d.get('roofline',{}).get('achieved'))
for k,v in d['step_model']['kernel_ms_by_kind'].items(): print(f"  {k:14s} {v}")
P
