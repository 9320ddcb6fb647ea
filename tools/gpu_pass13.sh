#!/bin/bash
# e2e spread after the GC freeze (3 runs) + ncu full capture with source counters of the six slow thin
# tapgemm launches of one step (1x1 K=32, ConvTranspose fwd / dgrad at 128x128)
mkdir -p gpurun_out
for r in 1 2 3; do
timeout 900 python bench.py --no-cpu-baseline --no-kernel-timing --steps 30 > gpurun_out/bench_e2e_$r.json 2> gpurun_out/bench_e2e_$r.err
python -c "
import json;d=json.load(open('gpurun_out/bench_e2e_$r.json'));print('run $r value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),round(d['e2e']['ms_per_step'],3))"
done
B="python bench.py --steps 1 --warmup 3 --no-kernel-timing --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:tapgemm_kernel --launch-skip 21 -c 6 -f -o /tmp/thin $B > gpurun_out/ncu_thin.log 2>&1; echo "thin rc=$?"
ncu -i /tmp/thin.ncu-rep --page raw --csv > gpurun_out/ncu_raw_thin.csv 2>/dev/null
ncu -i /tmp/thin.ncu-rep --page source --csv > gpurun_out/ncu_src_thin.csv 2>/dev/null
ls -la gpurun_out/ncu_raw_thin.csv gpurun_out/ncu_src_thin.csv
