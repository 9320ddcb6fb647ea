#!/bin/bash
mkdir -p gpurun_out
for cfg in "resnet34 1 64 2" "resnet34 0 64 2" "resnet50 0 64 2" "resnet34 1 96 1" "resnet50 1 64 2" "resnet34 1 128 4"; do
  tag=$(echo $cfg | tr ' ' '_')
  timeout 600 python tools/model_debug.py $cfg > gpurun_out/model_$tag.log 2>&1; echo "model $cfg rc=$?"
  grep -v Warning gpurun_out/model_$tag.log | tail -n 48 | cut -c1-230
done
