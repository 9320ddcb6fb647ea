#!/bin/bash
# 2-GPU data-parallel evidence (tight timeouts: a collective mismatch must not burn the budget)
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:2}"; }
show() { python -c "
import json,sys;d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1]);print(sys.argv[1],'value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'par',d['config']['parallelism'],'loss',d['final_loss'])" $1; }
timeout 150 bash -c "$(declare -f run); run 29612 --steps 20 --warmup 3 --mode dp --no-cpu-baseline" > gpurun_out/bench_n2_dp.json 2> gpurun_out/bench_n2_dp.err; echo "dp rc=$?"; tail -3 gpurun_out/bench_n2_dp.err; show gpurun_out/bench_n2_dp.json
timeout 200 bash -c "$(declare -f run); run 29613 --steps 5 --warmup 3 --mode dp --size 1024 --batch 32 --no-cpu-baseline" > gpurun_out/bench_n2_c5_dp.json 2> gpurun_out/bench_n2_c5_dp.err; echo "c5 dp rc=$?"; tail -3 gpurun_out/bench_n2_c5_dp.err; show gpurun_out/bench_n2_c5_dp.json
