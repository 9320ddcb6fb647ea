#!/bin/bash
# pack kernel with compile-time taps, maxpool_bwd batched loads, SE weight-gradient batched loads,
# bias column sums on the side branch: full gpu suite, bench, in-graph costs, op times
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_b64.err
python -c "
import json;d=json.load(open('gpurun_out/bench_b64.json'));print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'loss',d['final_loss'],'launches',d['launches_per_step'],'frac',d['roofline']['frac'])"
timeout 500 python tools/graph_costs.py 64 256 > gpurun_out/graph_costs_b64.txt 2>&1; grep -E "==|pack|maxpool|se_mlp|colsum|without" gpurun_out/graph_costs_b64.txt
timeout 600 python tools/op_times.py 64 256 > gpurun_out/op_times_b64.txt 2>&1; grep -E "pack|maxpool|se_mlp|colsum " gpurun_out/op_times_b64.txt | head -20
