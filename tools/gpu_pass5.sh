#!/bin/bash
mkdir -p gpurun_out
python tools/step_segments.py 64 256 > gpurun_out/step_segments_b64.txt 2>&1; echo "segments rc=$?"; cat gpurun_out/step_segments_b64.txt | tail -12
