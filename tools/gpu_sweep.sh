#!/bin/bash
# batch sweep at 256x256 (how the fixed per-launch costs amortise): 128 and 256 images per step
mkdir -p gpurun_out
for B in 128 256; do
timeout 300 python bench.py --batch $B --steps 10 --no-cpu-baseline > gpurun_out/bench_b$B.json 2> gpurun_out/bench_b$B.err; echo "b$B rc=$?"; tail -2 gpurun_out/bench_b$B.err
python -c "
import json;d=json.load(open('gpurun_out/bench_b$B.json'));print('b$B value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'frac',round(d['roofline']['frac'],4),'tflops',round(d['roofline']['achieved'],1))"
done
