#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py -q -x > gpurun_out/pytest_conv.log 2>&1; echo "conv rc=$?"; tail -3 gpurun_out/pytest_conv.log
timeout 600 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_b64.err
python -c "
import json;d=json.load(open('gpurun_out/bench_b64.json'));k=d['step_model']['kernel_ms_by_kind'];print('value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'loss',d['final_loss'],'frac',round(d['roofline']['frac'],4),{n:k[n]['ms'] for n in k if 'wgrad' in n})"
