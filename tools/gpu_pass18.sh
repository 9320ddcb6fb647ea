#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_conv_gpu.py -q -x > gpurun_out/pytest_conv.log 2>&1; echo "conv rc=$?"; tail -2 gpurun_out/pytest_conv.log
run() { local name=$1; shift
  env "$@" timeout 200 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64_$name.json 2> gpurun_out/bench_b64_$name.err; echo "bench $name rc=$?"; tail -2 gpurun_out/bench_b64_$name.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_b64_$name.json'));k=d['step_model']['kernel_ms_by_kind'];print('$name value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'loss',d['final_loss'],'frac',round(d['roofline']['frac'],4),'wgrad3x3',k['wgrad3x3s1']['ms'])"; }
run red1 A=1
run red1_halo64 B2SEG_WGRAD_HALO=3
run base B2SEG_WGRAD_RED1_MIN=100000
