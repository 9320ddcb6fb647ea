mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_model_gpu.py -q -x -k "fused or 256-4 or 128-4" -s 2>&1 | grep -E "^\[|passed|failed|Error" | tail
run() { local name=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu-baseline --no-extras --steps 30 > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; echo "bench $name rc=$?"; tail -2 gpurun_out/bench_$name.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_$name.json'));print('$name value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'launches',d['launches_per_step'])"; }
run bwd0 B2SEG_BN_FUSED_BWD_MB=0
run bwd6 B2SEG_BN_FUSED_BWD_MB=6
run bwd12 B2SEG_BN_FUSED_BWD_MB=12
run bwd24 B2SEG_BN_FUSED_BWD_MB=24
run bwd48 B2SEG_BN_FUSED_BWD_MB=48
run bwdall B2SEG_BN_FUSED_BWD_MB=100000
run fwd1_bwd12 B2SEG_BN_FUSED_FWD=1 B2SEG_BN_FUSED_BWD_MB=12
B2SEG_BN_FUSED_BWD_MB=100000 timeout 600 python tools/op_times.py 64 256 > gpurun_out/op_times_b64_bwdall.txt 2>&1; grep "ew25" gpurun_out/op_times_b64_bwdall.txt | head -60
