#!/bin/bash
# ncu evidence: launch list of the timed steps (cudaProfilerStart/Stop bracket them in bench.py),
# and one --set full capture of the conv kernels.
mkdir -p gpurun_out
ARGS="--steps 2 --warmup 3 --no-kernel-timing --no-cpu-baseline"
python bench.py $ARGS > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err || { echo "plain bench failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv python bench.py $ARGS > gpurun_out/ncu_launch.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:"${KREGEX:-tapgemm_kernel|wgrad_kernel}" \
    --launch-skip ${KSKIP:-0} -c ${KCOUNT:-12} -f -o gpurun_out/conv_full python bench.py --steps 1 --warmup 3 --no-kernel-timing --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out
# second capture: tail of the backward pass (enc1 dgrad / wgrad, the wide-M narrow-N layers)
ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:"tapgemm_kernel|wgrad_kernel" --launch-skip 150 -c 18 -f -o gpurun_out/conv_full_bwd \
    python bench.py --steps 1 --warmup 3 --no-kernel-timing --no-cpu-baseline > gpurun_out/ncu_full2.log 2>&1
echo "full capture 2 rc=$?"
