#!/bin/bash
# One parameterised GPU recipe (run under gpurun):  tools/gpu_check.sh [tests] [bench] [ops] [ncu] [hbm] [dp]
#   tests  pytest -m gpu (+ smoke)          bench  default bench.py + reference arm
#   ops    per-op / per-segment / in-graph timings   ncu   launch list + DRAM bytes per kernel of one step
#   hbm    tools/hbm_bench.py ([64,1,1024,1024] loss / metric / augment kernels)   warp / warpncu   warp kernel: both tile shapes, ncu --set full
# Outputs land in gpurun_out/ (scratch); copy what should be judged into profiles/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
for what in "$@"; do case $what in
tests)
  timeout 2400 python -m pytest tests -q -m gpu -x -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
  grep -E "^\[|worst|passed|failed|Error|error" gpurun_out/pytest_gpu.log | tail -40
  timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log;;
bench)
  timeout 1200 python bench.py > gpurun_out/bench_b64.json 2> gpurun_out/bench_b64.err; echo "bench rc=$?"; tail -5 gpurun_out/bench_b64.err
  python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_b64.json'))
print('value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value'], 1), d['e2e'].get('ms_per_step_blocks'),
      'loss', d['final_loss'], 'launches', d['launches_per_step'])
for k in ('parity_check', 'sustained', 'roofline', 'c1_batch4', 'dp_c5', 'cpu_baseline'):
    print(k, json.dumps(d.get(k))[:600])
print('loss_hbm', json.dumps({k: (v['us'], v['frac']) for k, v in d.get('loss_hbm', {}).get('kernels', {}).items()}))
PY
  timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_ref.json;;
ops)
  timeout 300 python tools/step_segments.py 64 256 > gpurun_out/step_segments_b64.txt 2>&1; grep -E "graph|full" gpurun_out/step_segments_b64.txt
  timeout 600 python tools/op_times.py 64 256 > gpurun_out/op_times_b64.txt 2>&1; echo "op_times rc=$?"
  timeout 300 python tools/graph_costs.py 64 256 > gpurun_out/graph_costs_b64.txt 2>&1; echo "graph_costs rc=$?"; grep -E "^==|without" gpurun_out/graph_costs_b64.txt;;
ncu)
  A="--steps 2 --warmup 3 --no-kernel-timing --no-cpu-baseline --no-extras"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
      --log-file gpurun_out/launches.csv python bench.py $A > gpurun_out/ncu_launch.log 2>&1; echo "launch list rc=$?"
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
      --profile-from-start off --csv --log-file gpurun_out/dram_step.csv \
      python bench.py --steps 1 --warmup 3 --no-kernel-timing --no-cpu-baseline --no-extras > gpurun_out/ncu_dram.log 2>&1; echo "dram list rc=$?"
  python tools/dram_summary.py gpurun_out/dram_step.csv --json gpurun_out/dram_step_families.json > gpurun_out/dram_step_summary.txt 2>&1; head -30 gpurun_out/dram_step_summary.txt;;
dp)
  # data parallel over all visible GPUs: consistency over real NCCL, then the bench line (carries dp_c5)
  NG=$(nvidia-smi -L | wc -l)
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29631 \
      tools/dp_check.py > gpurun_out/dp_check_n$NG.txt 2>&1; echo "dp_check rc=$?"; grep -E "OK|FAIL|PASSED|FAILED|Error" gpurun_out/dp_check_n$NG.txt | tail -20
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29632 \
      bench.py --gpus $NG > gpurun_out/bench_n$NG.json 2> gpurun_out/bench_n$NG.err; echo "bench N=$NG rc=$?"; tail -3 gpurun_out/bench_n$NG.err
  python - <<PY
import json
d = json.load(open('gpurun_out/bench_n$NG.json'))
print('value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'n', d['n_gpus'], 'e2e', round(d['e2e']['value'], 1))
print('dp_c5', json.dumps(d.get('dp_c5')))
PY
  NCCL_DEBUG=INFO timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29633 \
      bench.py --gpus $NG --steps 2 --warmup 3 --no-extras --mode dp > gpurun_out/bench_dp_n$NG.json 2> gpurun_out/nccl_info_n$NG.txt; echo "dp-mode bench rc=$?"
  grep -E "NVLS|via P2P|Channel 00/" gpurun_out/nccl_info_n$NG.txt | head -5; cut -c1-400 gpurun_out/bench_dp_n$NG.json;;
hbm)
  timeout 600 python tools/hbm_bench.py > gpurun_out/hbm_kernels.json 2> gpurun_out/hbm_kernels.err; echo "hbm rc=$?"; tail -3 gpurun_out/hbm_kernels.err;;
warp)
  # warp kernel, both tile shapes: bit-exactness against cv2, then [128 planes, 1024^2] timings
  for t in 32x32 64x16; do
    B2SEG_WARP_TILE=$t timeout 900 python -m pytest tests/test_augment_gpu.py tests/test_augment_ops_gpu.py -q -m gpu -x \
        > gpurun_out/pytest_warp_$t.log 2>&1; echo "warp tile $t pytest rc=$?"; tail -1 gpurun_out/pytest_warp_$t.log
    B2SEG_WARP_TILE=$t timeout 300 python tools/hbm_bench.py --only aug > gpurun_out/hbm_aug_$t.json 2> gpurun_out/hbm_aug_$t.err
    echo "hbm tile $t rc=$?"; grep -A1 "warp\|remap" gpurun_out/hbm_aug_$t.json | grep -o '"aug[a-z_]*"\|"us": [0-9.]*' | paste - - - - - -
  done;;
warpncu)
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:warp_u8 -s 4 -c 3 -f -o gpurun_out/warp \
      python tools/warp_profile.py > gpurun_out/warp_ncu.log 2>&1; echo "warp ncu rc=$?"; tail -2 gpurun_out/warp_ncu.log;;
bnab)
  # forward BatchNorm A/B: finalize kernels (0) against fixed-point accumulators + fused apply (2)
  for m in 0 2 0 2; do
    B2SEG_BN_FUSED_FWD=$m timeout 600 python bench.py --no-cpu-baseline --no-extras > gpurun_out/bench_bnfwd$m.json 2> gpurun_out/bench_bnfwd$m.err
    echo "BN_FUSED_FWD=$m rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_bnfwd$m.json'));print(round(d['ms_per_step'],3),'ms', d['launches_per_step'],'launches', d['final_loss'], d.get('parity_check'))"
  done;;
esac; done
ls gpurun_out | wc -l
