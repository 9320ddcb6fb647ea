#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py -q -x > gpurun_out/pytest_conv.log 2>&1; echo "conv rc=$?"; tail -3 gpurun_out/pytest_conv.log
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_b64.err
python -c "
import json;d=json.load(open('gpurun_out/bench_b64.json'));k=d['step_model']['kernel_ms_by_kind'];print('value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'loss',d['final_loss'],'frac',round(d['roofline']['frac'],4),{n:k[n]['ms'] for n in ('dgrad1x1s1','convT_dgrad','conv1x1s1','convT2x2','dgrad3x3s2')})"
timeout 600 python tools/op_times.py 64 256 > gpurun_out/op_times_b64.txt 2>&1; grep -E "dgrad1x1s1|convT_dgrad" gpurun_out/op_times_b64.txt
