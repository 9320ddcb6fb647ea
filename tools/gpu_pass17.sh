#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_b64.err
python -c "
import json;d=json.load(open('gpurun_out/bench_b64.json'));k=d['step_model']['kernel_ms_by_kind'];print('value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'loss',d['final_loss'],'maxpool_fwd', k.get('ew6'))"
