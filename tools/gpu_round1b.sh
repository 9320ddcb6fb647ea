#!/bin/bash
# GPU pass: full gpu test-suite, smoke, short bench (C1 batch 4 and throughput batch 64)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --batch 4 --no-cpu-baseline > gpurun_out/bench_b4.json 2> gpurun_out/bench_b4.err; echo "bench b4 rc=$?"; tail -2 gpurun_out/bench_b4.err; cat gpurun_out/bench_b4.json
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench b64 rc=$?"; tail -2 gpurun_out/bench_b64.err; cat gpurun_out/bench_b64.json
