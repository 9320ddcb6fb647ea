#!/bin/bash
# ncu evidence: launch list of two timed steps, and one --set full capture of the conv kernels.
mkdir -p gpurun_out
ARGS="--steps 2 --warmup 3 --no-kernel-timing --no-cpu-baseline"
python bench.py $ARGS > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip ${SKIP:-1800} -c ${COUNT:-1300} --csv \
    --log-file gpurun_out/launches.csv python bench.py $ARGS > gpurun_out/ncu_launch.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'tapgemm_kernel|wgrad_kernel' \
    --launch-skip ${KSKIP:-520} -c ${KCOUNT:-6} -f -o gpurun_out/conv_full python bench.py $ARGS > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out
