#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stem_gpu.py -q -x > gpurun_out/pytest_stem.log 2>&1; echo "stem rc=$?"; tail -12 gpurun_out/pytest_stem.log
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_b64.err
python -c "
import json;d=json.load(open('gpurun_out/bench_b64.json'));print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'loss',d['final_loss'],'launches',d['launches_per_step'])"
python tools/step_segments.py 64 256 2>&1 | grep -E "graph|full"
python tools/op_times.py 64 256 > gpurun_out/op_times_b64.txt 2>&1; grep -E "stem|maxpool|pack" gpurun_out/op_times_b64.txt
