#!/bin/bash
# augment tests + HBM bench at HEAD, per-op times, ncu --set full of the loss kernels and the conv kernels
# (reports converted to raw CSV on the box: gpurun_out/ is capped at 64 MiB)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_augment_gpu.py -q -x > gpurun_out/pytest_aug.log 2>&1; echo "pytest aug rc=$?"; tail -3 gpurun_out/pytest_aug.log
timeout 600 python tools/hbm_bench.py > gpurun_out/hbm_kernels.json 2> gpurun_out/hbm_kernels.err; echo "hbm rc=$?"
python -c "
import json;h=json.load(open('gpurun_out/hbm_kernels.json'))
for k,v in h['kernels'].items(): print(k, v['us'], v['gbs'], v['frac_of_peak'])"
python tools/op_times.py 64 256 > gpurun_out/op_times_b64.txt 2>&1; echo "op_times rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"seg_loss|seg_metrics" -c 5 -f -o /tmp/loss_full \
   python tools/hbm_bench.py --iters 1 > gpurun_out/ncu_loss.log 2>&1; echo "ncu loss rc=$?"
ncu -i /tmp/loss_full.ncu-rep --page raw --csv > gpurun_out/ncu_full_loss_raw.csv 2>/dev/null
timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:"halo3x3_kernel|wgrad_kernel|tapgemm_kernel" -c 60 -f -o /tmp/conv_full_head \
    python bench.py --steps 1 --warmup 3 --no-kernel-timing --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu conv rc=$?"
ncu -i /tmp/conv_full_head.ncu-rep --page raw --csv > gpurun_out/ncu_full_conv_raw.csv 2>/dev/null
ls -la gpurun_out /tmp/*.ncu-rep
