#!/bin/bash
# ncu --set full of the largest HBM-bound layer kernels of one step (raw CSV only: 64 MiB cap)
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --no-kernel-timing --no-cpu-baseline"
cap() { timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"$1" -c $2 -f -o /tmp/$3 $B > gpurun_out/ncu_$3.log 2>&1; echo "$3 rc=$?"; ncu -i /tmp/$3.ncu-rep --page raw --csv > gpurun_out/ncu_raw_$3.csv 2>/dev/null; }
cap "stem_fwd|maxpool|stem_wgrad" 4 stem
cap "bn_bwd_reduce|bn_bwd_apply|gate_bwd|dot1_bwd" 10 bnbwd
cap "bn_apply|pack_kernel" 4 bnfwd
ncu -i /tmp/stem.ncu-rep --page source --csv -k regex:maxpool_bwd > gpurun_out/ncu_src_maxpool_bwd.csv 2>/dev/null
ls -la gpurun_out | head -30
