#!/bin/bash
# GPU pass: gpu test-suite, smoke, default bench (with CPU baseline), reference arm, HBM kernel bench,
# ncu launch list of the timed steps.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench b64 rc=$?"; tail -2 gpurun_out/bench_b64.err; cut -c1-1500 gpurun_out/bench_b64.json
timeout 600 python tools/hbm_bench.py > gpurun_out/hbm_kernels.json 2> gpurun_out/hbm_kernels.err; echo "hbm rc=$?"; tail -3 gpurun_out/hbm_kernels.err; cat gpurun_out/hbm_kernels.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
ARGS="--steps 2 --warmup 3 --no-kernel-timing --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python bench.py $ARGS > gpurun_out/ncu_launch.log 2>&1
echo "launch list rc=$?"
