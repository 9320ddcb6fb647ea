#!/bin/bash
# 2-GPU check of the torchrun paths the driver's scaling run uses
mkdir -p gpurun_out
nvidia-smi -L | head -4
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:2}"; }
timeout 600 bash -c "$(declare -f run); run 29611 --steps 20 --warmup 3" > gpurun_out/bench_n2_folds.json 2> gpurun_out/bench_n2_folds.err; echo "folds rc=$?"; tail -3 gpurun_out/bench_n2_folds.err; cut -c1-400 gpurun_out/bench_n2_folds.json
timeout 600 bash -c "$(declare -f run); run 29612 --steps 20 --warmup 3 --mode dp" > gpurun_out/bench_n2_dp.json 2> gpurun_out/bench_n2_dp.err; echo "dp rc=$?"; tail -3 gpurun_out/bench_n2_dp.err; cut -c1-400 gpurun_out/bench_n2_dp.json
timeout 600 bash -c "$(declare -f run); run 29613 --steps 2 --warmup 1 --impl reference" > gpurun_out/bench_n2_ref.json 2> gpurun_out/bench_n2_ref.err; echo "ref rc=$?"; tail -3 gpurun_out/bench_n2_ref.err; cut -c1-300 gpurun_out/bench_n2_ref.json
