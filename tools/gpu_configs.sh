#!/bin/bash
# other configs at HEAD: C5 shape on one GPU (1024x1024, batch 64) and the resnet50 backbone of C4
mkdir -p gpurun_out
show() { python -c "
import json,sys;d=json.load(open(sys.argv[1]));print(sys.argv[1],'value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'frac',round(d['roofline']['frac'],4),'tflops',round(d['roofline']['achieved'],1),'loss',d['final_loss'])" $1; }
timeout 400 python bench.py --size 1024 --batch 64 --steps 5 --no-cpu-baseline > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; echo "c5 rc=$?"; tail -2 gpurun_out/bench_c5_n1.err; show gpurun_out/bench_c5_n1.json
timeout 400 python bench.py --backbone resnet50 --batch 64 --steps 10 --no-cpu-baseline > gpurun_out/bench_r50.json 2> gpurun_out/bench_r50.err; echo "r50 rc=$?"; tail -2 gpurun_out/bench_r50.err; show gpurun_out/bench_r50.json
timeout 300 python bench.py --no-attention --loss focal --steps 10 --no-cpu-baseline > gpurun_out/bench_noattn_focal.json 2> gpurun_out/bench_noattn_focal.err; echo "noattn rc=$?"; show gpurun_out/bench_noattn_focal.json
