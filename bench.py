#!/usr/bin/env python
"""Headline benchmark: U-Net training images/s at 256x256 on B200 (BASELINE.json `metric`).

    python bench.py --gpus N --steps K --warmup W            # the B200 path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)
    torchrun ... bench.py --gpus N ...                       # one rank per GPU

A step = one optimisation step of `UNetWithBackbone(backbone='resnet34', use_attention=True)`
(the constructor default minus the weight download) with ComboLoss (BCE+Dice) and Adam
(lr 1e-3, wd 1e-5) on a batch of synthetic 256x256 microscopy tiles: forward, loss + metrics,
backward, optimizer.step() -- the reference's inner loop, train.py:38-60.

Printed JSON (one line, rank 0):
  value        images/s, whole job, inputs resident in HBM, CUDA events, max over ranks
  e2e          same metric through train.train_step with HOST (pinned) batches fed by
               train.DevicePrefetcher (H2D on a side stream, double buffered) and the loss +
               confusion counts read back every step
  roofline     dominant kernel class (tcgen05 conv GEMMs): algorithmic FLOPs / event-timed duration
  cpu_baseline the oracle port of the reference step on this box's host cores (rank 0, N=1)
Multi-GPU (`--mode folds`, default): every rank trains its own model on its own data -- the
cross-validation-fold / grid-config sharding of north_star, no data-path collective (weak
scaling).  `--mode dp`: one model, per-rank batch, flat-gradient NCCL all-reduce per step.
"""
import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "image-segmentation-project_b200")
for _p in (PKG, ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "unet_train_images_per_sec_256x256"
UNIT = "images/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=int(os.environ.get("B2SEG_BENCH_BATCH", "64")),
                    help="images per GPU per step")
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--backbone", default="resnet34")
    ap.add_argument("--no-attention", action="store_true")
    ap.add_argument("--loss", default="combo")
    ap.add_argument("--mode", default="folds", choices=["folds", "dp"])
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-timing", action="store_true")
    ap.add_argument("--torch-adam", action="store_true", help="stock torch.optim.Adam instead of FusedAdam")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    return a


def workload(a):
    return {
        "workload": f"UNetWithBackbone({a.backbone}, use_attention={not a.no_attention}, pretrained=False) "
                    f"+ {a.loss} loss + Adam(lr=1e-3, wd=1e-5), synthetic {a.size}x{a.size} grayscale cell "
                    f"tiles + binary masks, {a.batch} images/GPU/step",
        "backbone": a.backbone, "use_attention": not a.no_attention, "loss_fn": a.loss,
        "img_size": a.size, "batch_per_gpu": a.batch,
    }


# ------------------------------------------------------------------ clocks ----
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
            self.thread.join(timeout=2)

    def summary(self):
        sm, smax, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(smax), "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------- B200 arm ----
def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return {"tensor": d.get("bf16_tflops_sustained", 1400.0), "tensor_burst": d.get("bf16_tflops"),
                "hbm": d.get("hbm_gbs"), "source": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    return {"tensor": 1400.0, "tensor_burst": 1590.0, "hbm": 6650.0,
            "source": "fallback (B200_PROFILING.md: 1.59 PF burst / ~1.4 PF sustained, 6.65 TB/s)"}


def _recorded_traffic(a):
    """DRAM bytes per step of the tcgen05 conv kernels from the committed ncu capture of this exact
    workload (profiles/r01_dram_step_families.json, made by tools/dram_summary.py from
    `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`); None for any other workload."""
    path = os.path.join(ROOT, "profiles", "r01_dram_step_families.json")
    if not (a.batch == 64 and a.size == 256 and a.backbone == "resnet34" and not a.no_attention
            and os.path.exists(path)):
        return None
    fam = json.load(open(path))["families"]
    return float(sum(fam[k]["rd"] + fam[k]["wr"] for k in ("halo3x3_kernel", "tapgemm_kernel", "wgrad_kernel")
                     if k in fam))


def _dist_setup(a):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    elif a.gpus > 1:
        print(f"bench.py: --gpus {a.gpus} needs torchrun (one rank per GPU); running 1 rank", file=sys.stderr)
    return world, rank, local


def _barrier(world):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def _max_over_ranks(ms, world, dev):
    import torch
    import torch.distributed as dist
    if world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _kernel_breakdown(model, crit, images, masks):
    """Event-timed replay of the planned forward / backward sequences (one extra step, outside
    the throughput region): per-op-kind time and the tensor-core ops' algorithmic FLOP rate."""
    import torch
    model.train()
    out = model(images)
    loss = crit(out, masks)
    plan = model._plan_for(images, True)
    torch.cuda.synchronize()
    agg = {}
    for rep in range(3):
        for seq in (plan.fwd, plan.bwd):
            ms = seq.run_timed()
            if rep == 0:
                continue  # cold replay
            for op, t in zip(seq.ops, ms):
                k = agg.setdefault(op.kind, [0.0, 0.0, 0])
                k[0] += t / 2
                k[1] += op.flops / 2
                k[2] += 1
    del loss, out
    tens_ms = sum(v[0] for v in agg.values() if v[1] > 0)
    tens_fl = sum(v[1] for v in agg.values() if v[1] > 0)
    n_tens = sum(v[2] for v in agg.values() if v[1] > 0) // 2
    total_ms = sum(v[0] for v in agg.values())
    kinds = {k: {"ms": round(v[0], 4), "launches": v[2] // 2,
                 **({"tflops": round(v[1] / v[0] / 1e9, 1)} if v[1] > 0 and v[0] > 0 else {})}
             for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])}
    return {"tensor_ms": tens_ms, "tensor_flops": tens_fl, "tensor_launches": n_tens,
            "total_ms": total_ms, "kinds": kinds}


def run_b200(a):
    import torch
    import torch.distributed as dist
    world, rank, local = _dist_setup(a)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: the B200 arm needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    import advanced_models
    import losses
    import train as train_mod
    from b2seg import synth

    torch.manual_seed(1234 + (rank if a.mode == "folds" else 0))
    model = advanced_models.UNetWithBackbone(1, a.backbone, pretrained=False,
                                             use_attention=not a.no_attention).to(dev)
    crit = losses.get_loss_function({"loss_fn": a.loss})
    from b2seg.optim import FusedAdam, LAUNCHES_PER_GROUP
    # the optimizer train.quick_train / cross_validation build: a torch.optim.Adam subclass whose
    # step() is one multi-tensor kernel launch (--torch-adam times the stock optimizer instead)
    opt = (torch.optim.Adam if a.torch_adam else FusedAdam)(model.parameters(), lr=1e-3, weight_decay=1e-5)
    if a.mode == "dp" and world > 1:
        from b2seg import dataparallel
        dataparallel.attach(model, world, criterion=crit)  # global-batch loss, summed gradients
    model.train()

    # synthetic pool: POOL distinct batches, pinned on the host and resident on the device
    POOL = 4
    tiles = min(a.batch * POOL, 64)
    base_i, base_m = synth.cell_tiles(tiles, a.size, seed=1234 + rank)
    idx = torch.arange(a.batch * POOL) % tiles
    host_i = base_i[idx].view(POOL, a.batch, 1, a.size, a.size).contiguous().pin_memory()
    host_m = base_m[idx].view(POOL, a.batch, 1, a.size, a.size).contiguous().pin_memory()
    dev_i, dev_m = host_i.to(dev), host_m.to(dev)

    def step_resident(i):
        return train_mod.train_step(model, dev_i[i % POOL], dev_m[i % POOL], opt, crit)

    # end to end = the loop train.train_epoch runs: host (pinned) batches through the
    # DevicePrefetcher (H2D of batch k+1 on a side stream while step k computes), train_step, and a
    # device->host read of every step's loss + confusion counts into pinned memory.
    res_loss = torch.empty(8, dtype=torch.float32).pin_memory()
    res_stats = torch.empty(8, 16, dtype=torch.float64).pin_memory()

    def host_batches(n):
        for i in range(n):
            yield host_i[i % POOL], host_m[i % POOL]

    def run_e2e(n):
        for i, (im, mk) in enumerate(train_mod.DevicePrefetcher(host_batches(n), dev)):
            loss, stats = train_mod.train_step(model, im, mk, opt, crit)
            res_loss[i % 8].copy_(loss, non_blocking=True)   # D2H of the step's result
            res_stats[i % 8].copy_(stats, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(res_loss[(n - 1) % 8])

    for i in range(a.warmup):
        step_resident(i)
    run_e2e(2)
    # the plan holds ~10^5 long-lived Python objects: keep the cyclic collector from walking them
    # inside a timed region (a full collection costs tens of ms, several steps' worth)
    gc.collect()
    gc.freeze()
    _barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        torch.cuda.profiler.start()  # ncu --profile-from-start off captures exactly the timed steps
        e0.record()
        for i in range(a.steps):
            loss, _ = step_resident(i)
        e1.record()
        torch.cuda.profiler.stop()
        _barrier(world)
    ms = _max_over_ranks(e0.elapsed_time(e1), world, dev)
    final_loss = float(loss)

    # end to end: host batches in, loss/metrics out, every step
    run_e2e(2)
    _barrier(world)
    e0.record()
    run_e2e(a.steps)
    e1.record()
    _barrier(world)
    ms_e2e = _max_over_ranks(e0.elapsed_time(e1), world, dev)

    plan = model._plan_for(dev_i[0], True)
    from b2seg import lossfn
    launches_step = (plan.launches_fwd + plan.launches_bwd + lossfn.LAUNCHES_FWD + lossfn.LAUNCHES_BWD +
                     (0 if a.torch_adam else LAUNCHES_PER_GROUP))
    imgs = a.batch * world * a.steps
    peaks = _peaks()
    rec = {
        "metric": METRIC, "value": imgs / (ms / 1e3), "unit": UNIT, "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": dict(workload(a), global_batch=a.batch * world,
                       parallelism=("folds" if a.mode == "folds" else "dp") + str(world),
                       l2_policy=f"no explicit flush: one step streams {plan.nbytes / 2**20:.0f} MiB of "
                                 "activations/gradients (> 126 MB L2) and rotates over 4 input batches"),
        "e2e": {"value": imgs / (ms_e2e / 1e3), "unit": UNIT, "ms_per_step": ms_e2e / a.steps,
                "h2d_bytes_per_step": int(host_i[0].numel() * 4 + host_m[0].numel() * 4),
                "d2h_bytes_per_step": 4 + 16 * 8},
        "gpu_launches": launches_step * a.steps,
        "launches_per_step": launches_step,
        "clocks": clk.summary(),
        "final_loss": final_loss,
    }
    if a.mode == "dp" and world > 1:
        # the per-kernel replay below runs on rank 0 alone: take the collectives (gradient hook,
        # global-batch loss sums) off the model first, on every rank
        dataparallel.detach(model, crit)
    if rank == 0 and not a.no_kernel_timing:
        kb = _kernel_breakdown(model, crit, dev_i[0], dev_m[0])
        ach = kb["tensor_flops"] / (kb["tensor_ms"] * 1e-3) / 1e12
        rec["roofline"] = {
            "bound": "tensor", "kernel": "tapgemm_kernel / wgrad_kernel (tcgen05 conv fprop, dgrad, wgrad)",
            "achieved": ach, "peak": peaks["tensor"], "unit": "TFLOP/s", "frac": ach / peaks["tensor"],
            "traffic": _recorded_traffic(a), "traffic_unit": "DRAM bytes per step over these launches (ncu)",
            "peak_source": peaks["source"],
            "launches_per_step": kb["tensor_launches"],
            "flops_per_step": kb["tensor_flops"], "ms_per_step": kb["tensor_ms"],
            "share_of_step": kb["tensor_ms"] / kb["total_ms"],
        }
        rec["step_model"] = {
            "algorithmic_gflop_per_image": kb["tensor_flops"] / a.batch / 1e9,
            "whole_step_tflops": kb["tensor_flops"] / (ms / a.steps * 1e-3) / 1e12,
            "whole_step_frac_of_peak": kb["tensor_flops"] / (ms / a.steps * 1e-3) / 1e12 / peaks["tensor"],
            "kernel_ms_by_kind": kb["kinds"],
        }
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        rec["cpu_baseline"] = cpu_baseline(a, a.cpu_seconds)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        _emit(rec)


# ------------------------------------------------ CPU baseline / reference arm ----
def _oracle_trainer(a):
    """The reference step on the host cores through the oracle port (oracle/train_step.py):
    the same ATen fp32 CPU kernels the reference's nn.Modules reach.  /root/reference is a
    Python tree that cannot travel to the GPU box, hence kind = "port"."""
    import torch
    from oracle.state_layout import reference_state_shapes
    from oracle.train_step import OracleTrainer
    from oracle import unet_functional as uf
    shapes = reference_state_shapes(a.backbone, not a.no_attention)
    sd = uf.synth_state(shapes, seed=1)
    for k in sd:  # fresh-module BatchNorm state
        if k.endswith("running_mean"):
            sd[k].zero_()
        elif k.endswith("running_var"):
            sd[k].fill_(1.0)
    return OracleTrainer(sd, a.backbone, not a.no_attention, a.loss)


def _cpu_steps(a, batch, steps, warmup, budget_s=None):
    """Times `steps` oracle steps on batches of `batch` images; returns (seconds, steps done)."""
    import torch
    from b2seg import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    trainer = _oracle_trainer(a)
    imgs, masks = synth.cell_tiles(batch, a.size, seed=1234)
    for _ in range(warmup):
        trainer.step(imgs, masks)
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        trainer.step(imgs, masks)
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s and done >= 2:
            break
    return time.perf_counter() - t0, done, cores


def cpu_baseline(a, budget_s):
    batch = min(a.batch, 8)
    sec, done, cores = _cpu_steps(a, batch, steps=10 ** 6, warmup=1, budget_s=budget_s)
    return {"value": batch * done / sec, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{done} steps of {batch} images ({a.size}x{a.size}) of the same workload after 1 warm-up "
                      f"step, fp32, torch CPU with {cores} threads; oracle/train_step.py (forward, loss, "
                      "backward, Adam, metrics)",
            "ms_per_step": sec / done * 1e3}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return  # one CPU arm per box: the host cores are shared
    batch = min(a.batch, 8)
    steps = max(1, a.steps)
    sec, done, cores = _cpu_steps(a, batch, steps=steps, warmup=max(0, a.warmup), budget_s=240.0)
    v = batch * done / sec
    sample = (f"each step = {batch} images ({a.size}x{a.size}), a bounded sample of the "
              f"{a.batch}-image batch; {done} timed steps; fp32 torch CPU, {cores} threads")
    rec = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": max(world, a.gpus),
        "steps": done, "warmup": a.warmup, "ms_per_step": sec / done * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(workload(a), global_batch=a.batch * max(world, a.gpus), parallelism="cpu"),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(rec)


_RESULT_OUT = None


def _claim_stdout():
    """stdout carries exactly ONE JSON line.  Libraries write there too (NCCL prints its version
    banner on stdout at init), so the real stdout is set aside for the result and file
    descriptor 1 is pointed at stderr for everything else."""
    global _RESULT_OUT
    sys.stdout.flush()
    _RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def _emit(rec):
    out = _RESULT_OUT or sys.stdout
    out.write(json.dumps(rec) + "\n")
    out.flush()


def main():
    a = parse_args()
    _claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
