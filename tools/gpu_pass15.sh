#!/bin/bash
# data gradient enqueued before the weight gradient + side branch at low stream priority: A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
run() { local name=$1; shift
  env "$@" timeout 600 python bench.py --no-cpu-baseline --no-kernel-timing --steps 30 > gpurun_out/bench_b64_$name.json 2> gpurun_out/bench_b64_$name.err; echo "bench $name rc=$?"; tail -2 gpurun_out/bench_b64_$name.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_b64_$name.json'));print('$name value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'loss',d['final_loss'])"; }
run new A=1
run oldorder B2SEG_WGRAD_AFTER=0
run noprio B2SEG_SIDE_PRIO=0
run old_noprio B2SEG_WGRAD_AFTER=0 B2SEG_SIDE_PRIO=0
timeout 300 python tools/graph_costs.py 64 256 > gpurun_out/graph_costs_b64.txt 2>&1; grep -E "==|wgrad|without" gpurun_out/graph_costs_b64.txt
