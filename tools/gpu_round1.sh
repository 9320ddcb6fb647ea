#!/bin/bash
# first GPU pass: loss / augment parity, conv kernel error report
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 600 python -m pytest tests/test_loss_gpu.py tests/test_augment_gpu.py -q -m gpu > gpurun_out/pytest_loss_aug.log 2>&1; echo "loss/aug rc=$?"
tail -5 gpurun_out/pytest_loss_aug.log
for st in fprop dgrad wgrad; do
  timeout 300 python tools/conv_debug.py $st > gpurun_out/conv_$st.log 2>&1; echo "conv $st rc=$?"
done
tail -30 gpurun_out/conv_fprop.log
