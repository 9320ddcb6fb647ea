#!/bin/bash
# Round-end evidence at HEAD: gpu tests, smoke, default bench (+cpu baseline), reference arm, HBM kernel
# bench, per-op and per-segment timings, ncu launch list, per-kernel DRAM traffic of one step, and a
# --set full capture of the stem / BatchNorm / loss kernels (raw CSV only: gpurun_out is capped at 64 MiB).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench b64 rc=$?"; tail -2 gpurun_out/bench_b64.err
python -c "
import json;d=json.load(open('gpurun_out/bench_b64.json'));print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'loss',d['final_loss'],'launches',d['launches_per_step'],'cpu',d.get('cpu_baseline'))"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/bench_ref.json
timeout 600 python tools/hbm_bench.py > gpurun_out/hbm_kernels.json 2> gpurun_out/hbm_kernels.err; echo "hbm rc=$?"; tail -3 gpurun_out/hbm_kernels.err
timeout 300 python tools/step_segments.py 64 256 > gpurun_out/step_segments_b64.txt 2>&1; grep -E "graph|full" gpurun_out/step_segments_b64.txt
timeout 600 python tools/op_times.py 64 256 > gpurun_out/op_times_b64.txt 2>&1; echo "op_times rc=$?"
timeout 300 python tools/graph_costs.py 64 256 > gpurun_out/graph_costs_b64.txt 2>&1; echo "graph_costs rc=$?"
ARGS="--steps 2 --warmup 3 --no-kernel-timing --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python bench.py $ARGS > gpurun_out/ncu_launch.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --profile-from-start off --csv --log-file gpurun_out/dram_step.csv \
    python bench.py --steps 1 --warmup 3 --no-kernel-timing --no-cpu-baseline > gpurun_out/ncu_dram.log 2>&1
echo "dram list rc=$?"
B="python bench.py --steps 1 --warmup 3 --no-kernel-timing --no-cpu-baseline"
cap() { timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"$1" -c $2 -f -o /tmp/$3 $B > gpurun_out/ncu_$3.log 2>&1; echo "$3 rc=$?"; ncu -i /tmp/$3.ncu-rep --page raw --csv > gpurun_out/ncu_raw_$3.csv 2>/dev/null; }
cap "stem_tc|maxpool" 4 stemtc
cap "bn_bwd_reduce|bn_bwd_apply|bn_apply" 9 bn
cap "halo3x3_kernel|wgrad_halo_kernel|wgrad_kernel" 24 conv
ls -la gpurun_out | head -60
