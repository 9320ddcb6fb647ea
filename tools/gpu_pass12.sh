#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py -q -x > gpurun_out/pytest_conv.log 2>&1; echo "conv rc=$?"; tail -5 gpurun_out/pytest_conv.log
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
run() { # name, env...
  local name=$1; shift
  env "$@" timeout 900 python bench.py --no-cpu-baseline --steps 30 > gpurun_out/bench_b64_$name.json 2> gpurun_out/bench_b64_$name.err; echo "bench $name rc=$?"; tail -2 gpurun_out/bench_b64_$name.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_b64_$name.json'));k=d['step_model']['kernel_ms_by_kind'];print('$name value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'loss',d['final_loss'],'frac',round(d['roofline']['frac'],4),'wgrad3x3',k['wgrad3x3s1']['ms'],'conv3x3',k['conv3x3s1']['ms'],'dgrad3x3',k['dgrad3x3s1']['ms'])"
}
run default A=1
run bn128_96 B2SEG_TG_BN128_MIN=96
run halo0 B2SEG_WGRAD_HALO=0
run default2 A=1
timeout 600 python tools/op_times.py 64 256 > gpurun_out/op_times_b64.txt 2>&1
