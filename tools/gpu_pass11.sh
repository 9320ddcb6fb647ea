#!/bin/bash
# halo-reuse weight gradient: conv unit tests, full gpu suite, bench A/B, op times, in-graph costs
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py -q -x > gpurun_out/pytest_conv.log 2>&1; echo "conv rc=$?"; tail -15 gpurun_out/pytest_conv.log
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for H in 1 0; do
B2SEG_WGRAD_HALO=$H timeout 900 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64_halo$H.json 2> gpurun_out/bench_b64_halo$H.err; echo "bench halo=$H rc=$?"; tail -2 gpurun_out/bench_b64_halo$H.err
python -c "
import json;d=json.load(open('gpurun_out/bench_b64_halo$H.json'));print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'loss',d['final_loss'],'frac',d['roofline']['frac'],{k:v for k,v in d['step_model']['kernel_ms_by_kind'].items() if 'wgrad' in k})"
done
timeout 600 python tools/op_times.py 64 256 > gpurun_out/op_times_b64.txt 2>&1; grep -E "wgrad3x3s1" gpurun_out/op_times_b64.txt | sort -k4 -n -r | awk 'NR%4==1' | head -14
timeout 500 python tools/graph_costs.py 64 256 > gpurun_out/graph_costs_b64.txt 2>&1; grep -E "==|wgrad|without" gpurun_out/graph_costs_b64.txt
