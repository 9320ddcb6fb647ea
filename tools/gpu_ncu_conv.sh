#!/bin/bash
# ncu --set full with source on single conv layers (tools/conv_bench.py): tools/gpu_ncu_conv.sh <name> <kernel regex> <count> case...
mkdir -p gpurun_out
NAME=$1; RX=$2; CNT=$3; shift 3
REPS=1 python tools/conv_bench.py "$@" > gpurun_out/${NAME}_plain.log 2>&1 &&
REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:$RX -c $CNT \
    -o gpurun_out/$NAME python tools/conv_bench.py "$@" > gpurun_out/${NAME}_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${NAME}_ncu.log; ls -la gpurun_out/$NAME.ncu-rep
