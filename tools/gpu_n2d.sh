#!/bin/bash
# stdout of a 2-rank run must be exactly one JSON line (NCCL banner silenced); reference arm under torchrun
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:2}"; }
timeout 150 bash -c "$(declare -f run); run 29621 --steps 10 --warmup 3" > gpurun_out/n2_stdout.txt 2> gpurun_out/n2_stderr.txt; echo "folds rc=$?"
wc -l gpurun_out/n2_stdout.txt; cut -c1-200 gpurun_out/n2_stdout.txt
timeout 150 bash -c "$(declare -f run); run 29622 --steps 2 --warmup 1 --impl reference" > gpurun_out/n2_ref_stdout.txt 2> gpurun_out/n2_ref_stderr.txt; echo "ref rc=$?"
wc -l gpurun_out/n2_ref_stdout.txt; cut -c1-200 gpurun_out/n2_ref_stdout.txt
timeout 120 python bench.py --steps 5 --no-cpu-baseline > gpurun_out/n1_stdout.txt 2> gpurun_out/n1_stderr.txt; echo "n1 rc=$?"; wc -l gpurun_out/n1_stdout.txt; cut -c1-160 gpurun_out/n1_stdout.txt; tail -2 gpurun_out/n1_stderr.txt
