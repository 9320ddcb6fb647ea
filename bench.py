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
  cpu_baseline the reference's own train.train_epoch (oracle/_ref, staged by oracle/make_ref.py;
               the oracle port when it is absent) on this box's host cores (rank 0, N=1)
  parity_check the last timed step's loss and confusion counts against the CPU oracle evaluated
               on that step's stored logits
  sustained    the same resident step repeated for >= 3 s (clocks sampled), so the roofline
               fraction is also quoted at the clocks the sustained peak was measured at
  c1_batch4    BASELINE config 1's real shape (batch 4 x 256^2) and the scripts' (4 x 128^2)
  batch_sweep  the headline step at 128 and 256 images per step (N = 1 only)
  loss_hbm     loss forward / backward / metrics kernels at [64,1,1024,1024]: us, GB/s, frac of HBM peak
  dp_c5        config 5: 1024^2 tiles, global batch 64 split over the ranks, data parallel with the
               NCCL gradient all-reduce (at N = 1 the same step without a collective)
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
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the sub-records (sustained, c1_batch4, loss_hbm, dp_c5)")
    ap.add_argument("--sustained-seconds", type=float, default=3.0)
    ap.add_argument("--e2e-steps", type=int, default=200, help="minimum steps of the end-to-end measurement")
    ap.add_argument("--c5-size", type=int, default=1024)
    ap.add_argument("--c5-global-batch", type=int, default=64)
    ap.add_argument("--c5-steps", type=int, default=8)
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
    workload (profiles/rNN_dram_step_families.json, made by tools/dram_summary.py from
    `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`); None for any other workload."""
    path = os.path.join(ROOT, "profiles", "r02_dram_step_families.json")  # the newest committed capture
    if not os.path.exists(path):
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


def _timed_steps(fn, n, world):
    """n calls of fn(i) between two events on the current stream, barrier + synchronize on both
    sides; returns elapsed ms (this rank)."""
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    _barrier(world)
    return e0.elapsed_time(e1)


def _parity_check(plan, crit, loss_name, loss, masks):
    """The step the bench just timed against the CPU oracle: loss_name evaluated in fp64 on the
    plan's stored logits of that step (oracle/seg_losses.py, pinned to the reference's losses.py by
    tests/golden) and the exact confusion counts.  Checker only: outside every timed region."""
    import torch
    from oracle import seg_losses
    from b2seg import lossfn
    logits = plan.logits.detach().float().cpu()
    m = masks.detach().float().cpu()
    ref = float(seg_losses.loss_value(loss_name, logits, m))
    ours = float(loss)
    out = {"loss": ours, "oracle_loss_on_same_logits": ref, "rel": abs(ours - ref) / max(abs(ref), 1e-30),
           "tolerance": 1e-3}
    st = getattr(crit, "last_stats", None)
    if st is not None:
        got = lossfn.metrics_dict(st.cpu())
        want = seg_losses.confusion_metrics(torch.sigmoid(logits), m)
        out["metrics_max_abs_diff"] = max(abs(got[k] - want[k]) for k in want)
    out["ok"] = bool(out["rel"] <= 1e-3 and out.get("metrics_max_abs_diff", 0.0) <= 1e-3)
    return out


def _loss_hbm(dev, peaks):
    """Loss / metric kernels at the loss-roofline shape of SURVEY 8(d), [64,1,1024,1024] fp32:
    8 B/px forward and metrics, 12 B/px backward; two buffer sets (537 MB per pass > L2)."""
    import torch
    import losses
    from b2seg import lossfn
    n, size, iters = 64, 1024, 10
    shape = (n, 1, size, size)
    npx = n * size * size
    g = torch.Generator(device="cpu").manual_seed(1234)
    xs = [(torch.randn((npx // 64,), generator=g) * 2).repeat(64).view(shape).to(dev) for _ in range(2)]
    ts = [(torch.rand((npx // 64,), generator=g) > 0.9).float().repeat(64).view(shape).to(dev) for _ in range(2)]
    hbm = peaks["hbm"] or 6650.0
    out = {"shape": list(shape), "hbm_peak_gbs": hbm, "kernels": {}}

    def timed(fn):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    def rec(name, ms, bpp):
        gbs = bpp * npx / (ms * 1e-3) / 1e9
        out["kernels"][name] = {"us": round(ms * 1e3, 2), "bytes_per_px": bpp, "gbs": round(gbs, 1),
                                "frac": round(gbs / hbm, 4)}

    with ClockSampler(dev.index or 0) as clk:
        for name in ("combo", "focal", "tversky"):
            spec = losses.get_loss_function({"loss_fn": name})._spec()
            saved = {}

            def fwd(i, spec=spec):
                saved["r"] = lossfn.loss_forward(xs[i & 1], ts[i & 1], spec)

            rec(f"seg_loss_fwd[{name}]", timed(fwd), 8)
            _, _, coef, sv = saved["r"]
            rec(f"seg_loss_bwd[{name}]", timed(lambda i, coef=coef, sv=sv: lossfn.loss_backward(sv, coef, None, shape)), 12)
            del saved, sv
        ps = [torch.sigmoid(x) for x in xs]
        rec("seg_metrics_from_probs", timed(lambda i: lossfn.metrics_from_probs(ps[i & 1], ts[i & 1])), 8)
    out["clocks"] = clk.summary()
    out["min_frac"] = min(k["frac"] for k in out["kernels"].values())
    del xs, ts, ps
    torch.cuda.empty_cache()
    return out


def _small_batch(a, dev, batch, size, steps=50):
    """Config 1's real shape: a fresh model at `batch` x `size`^2 (resident inputs, same step)."""
    import torch
    import advanced_models
    import losses
    import train as train_mod
    from b2seg import synth
    from b2seg.optim import FusedAdam
    torch.manual_seed(4321)
    model = advanced_models.UNetWithBackbone(1, a.backbone, pretrained=False,
                                             use_attention=not a.no_attention).to(dev).train()
    crit = losses.get_loss_function({"loss_fn": a.loss})
    opt = FusedAdam(model.parameters(), lr=1e-3, weight_decay=1e-5)
    im, mk = synth.cell_tiles(batch * 2, size, seed=99)
    im, mk = im.to(dev).view(2, batch, 1, size, size), mk.to(dev).view(2, batch, 1, size, size)
    t0 = time.perf_counter()
    train_mod.train_step(model, im[0], mk[0], opt, crit)
    torch.cuda.synchronize()
    first = time.perf_counter() - t0   # plan build + graph capture + first step (host wall clock)
    for i in range(4):
        train_mod.train_step(model, im[i & 1], mk[i & 1], opt, crit)
    with ClockSampler(dev.index or 0) as clk:
        ms = _timed_steps(lambda i: train_mod.train_step(model, im[i & 1], mk[i & 1], opt, crit), steps, 1)
    plan = model._plan_for(im[0], True)
    out = {"batch": batch, "img_size": size, "steps": steps, "ms_per_step": ms / steps,
           "images_per_sec": batch * steps / (ms / 1e3), "first_step_s": round(first, 3),
           "launches_per_step": plan.launches_fwd + plan.launches_bwd, "clocks": clk.summary()}
    del model, opt, plan
    torch.cuda.empty_cache()
    return out


def _dp_c5(a, dev, world, rank, peaks):
    """Config 5: 1024^2 tiles, global batch 64, bf16, one model replicated over the ranks, NCCL
    gradient all-reduce (b2seg.dataparallel).  Strong scaling: per-rank batch = 64 / N.  Reports
    tiles/s (max over ranks), and the same step with the collectives taken off (what they cost)."""
    import torch
    import advanced_models
    import losses
    import train as train_mod
    from b2seg import synth, dataparallel
    from b2seg.optim import FusedAdam
    if a.c5_global_batch % world:
        return {"skipped": f"global batch {a.c5_global_batch} does not divide over {world} ranks"}
    per = a.c5_global_batch // world
    size = a.c5_size
    torch.manual_seed(777)
    model = advanced_models.UNetWithBackbone(1, "resnet34", pretrained=False, use_attention=False).to(dev).train()
    crit = losses.get_loss_function({"loss_fn": "combo"})
    opt = FusedAdam(model.parameters(), lr=1e-3, weight_decay=1e-5)
    base_i, base_m = synth.cell_tiles(min(per, 4), size, seed=500 + rank)
    idx = torch.arange(per) % base_i.shape[0]
    im, mk = base_i[idx].to(dev), base_m[idx].to(dev)
    if world > 1:
        dataparallel.attach(model, world, criterion=crit)

    def step(i):
        train_mod.train_step(model, im, mk, opt, crit)

    for i in range(3):
        step(i)
    with ClockSampler(dev.index or 0) as clk:
        ms = _max_over_ranks(_timed_steps(step, a.c5_steps, world), world, dev)
    plan = model._plan_for(im, True)
    flops = sum(op.flops for seq in (plan.fwd, plan.bwd) for op in seq.ops) * world
    out = {"workload": f"UNetWithBackbone(resnet34, use_attention=False) + combo + Adam, {size}x{size} tiles, "
                       f"global batch {a.c5_global_batch} = {per}/GPU x {world}, data parallel",
           "n_gpus": world, "per_gpu_batch": per, "steps": a.c5_steps, "ms_per_step": ms / a.c5_steps,
           "tiles_per_sec": a.c5_global_batch * a.c5_steps / (ms / 1e3), "scaling": "strong",
           "whole_step_tflops": flops / (ms / a.c5_steps * 1e-3) / 1e12,
           "whole_step_frac_of_peak": flops / (ms / a.c5_steps * 1e-3) / 1e12 / (peaks["tensor"] * world),
           "clocks": clk.summary()}
    if world > 1:
        out["allreduce_bytes"] = int(plan.gflat.numel() * 4)
        out["buckets"] = getattr(model, "_dp_buckets", 1)
        out["bucket_bytes"] = [int((hi - lo) * 4) for _, lo, hi in plan.buckets]
        dataparallel.detach(model, crit)
        os.environ["B2SEG_DP_BUCKETS"] = "0"   # A/B: one all-reduce after the whole backward
        dataparallel.attach(model, world, criterion=crit)
        for i in range(2):
            step(i)
        ms1 = _max_over_ranks(_timed_steps(step, a.c5_steps, world), world, dev)
        out["ms_per_step_single_allreduce"] = ms1 / a.c5_steps
        os.environ.pop("B2SEG_DP_BUCKETS", None)
        dataparallel.detach(model, crit)
        for i in range(2):
            step(i)
        ms0 = _max_over_ranks(_timed_steps(step, a.c5_steps, world), world, dev)
        out["ms_per_step_without_collectives"] = ms0 / a.c5_steps
        out["exposed_collective_ms"] = (ms - ms0) / a.c5_steps
    del model, opt, plan
    torch.cuda.empty_cache()
    return out


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

    def run_e2e(n, marks=None, every=0):
        """n end-to-end steps; with `marks`, a CUDA event is recorded every `every` steps."""
        for i, (im, mk) in enumerate(train_mod.DevicePrefetcher(host_batches(n), dev)):
            loss, stats = train_mod.train_step(model, im, mk, opt, crit)
            res_loss[i % 8].copy_(loss, non_blocking=True)   # D2H of the step's result
            res_stats[i % 8].copy_(stats, non_blocking=True)
            if marks is not None and (i + 1) % every == 0:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                marks.append(ev)
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
    plan = model._plan_for(dev_i[0], True)
    parity = None
    if rank == 0 and not (a.mode == "dp" and world > 1):
        parity = _parity_check(plan, crit, a.loss, loss, dev_m[(a.steps - 1) % POOL])

    # end to end: host batches in, loss/metrics out, every step.  Measured over >= e2e_steps steps
    # in blocks of `a.steps` (an event per block) so the spread is visible next to the mean.
    blocks = max(1, -(-max(a.e2e_steps, a.steps) // a.steps))
    n_e2e = blocks * a.steps
    run_e2e(2)
    _barrier(world)
    marks = []
    e0.record()
    run_e2e(n_e2e, marks, a.steps)
    e1.record()
    _barrier(world)
    ms_e2e = _max_over_ranks(e0.elapsed_time(e1), world, dev)
    prev, block_ms = e0, []
    for ev in marks:
        block_ms.append(prev.elapsed_time(ev) / a.steps)
        prev = ev
    block_ms.sort()

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
        "e2e": {"value": a.batch * world * n_e2e / (ms_e2e / 1e3), "unit": UNIT, "ms_per_step": ms_e2e / n_e2e,
                "steps": n_e2e,
                "ms_per_step_blocks": {"block_steps": a.steps, "min": block_ms[0],
                                       "median": block_ms[len(block_ms) // 2], "max": block_ms[-1]},
                "h2d_bytes_per_step": int(host_i[0].numel() * 4 + host_m[0].numel() * 4),
                "d2h_bytes_per_step": 4 + 16 * 8},
        "gpu_launches": launches_step * a.steps,
        "launches_per_step": launches_step,
        "clocks": clk.summary(),
        "final_loss": final_loss,
    }
    if parity is not None:
        rec["parity_check"] = parity
    extras = not a.no_extras
    if extras:
        # the same resident step for >= sustained_seconds: boost clocks decay to what the chip holds
        n_sus = max(a.steps, int(a.sustained_seconds * 1e3 / (ms / a.steps)) + 1)
        with ClockSampler(local) as clk_s:
            ms_s = _max_over_ranks(_timed_steps(step_resident, n_sus, world), world, dev)
        rec["sustained"] = {"steps": n_sus, "seconds": ms_s / 1e3, "ms_per_step": ms_s / n_sus,
                            "value": a.batch * world * n_sus / (ms_s / 1e3), "unit": UNIT,
                            "clocks": clk_s.summary()}
    if a.mode == "dp" and world > 1:
        # the per-kernel replay below runs on rank 0 alone: take the collectives (gradient hook,
        # global-batch loss sums) off the model first, on every rank
        dataparallel.detach(model, crit)
    if rank == 0 and not a.no_kernel_timing:
        kb = _kernel_breakdown(model, crit, dev_i[0], dev_m[0])
        ach = kb["tensor_flops"] / (kb["tensor_ms"] * 1e-3) / 1e12
        rec["roofline"] = {
            "bound": "tensor", "kernel": "halo3x3_kernel / tapgemm_kernel / wgrad_kernel (tcgen05 conv fprop, dgrad, wgrad)",
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
        if "sustained" in rec:
            rec["sustained"]["whole_step_frac_of_peak"] = (
                kb["tensor_flops"] / (rec["sustained"]["ms_per_step"] * 1e-3) / 1e12 / peaks["tensor"])
    # free the headline model before the sub-records build theirs
    del plan
    model._plans.clear()
    torch.cuda.empty_cache()
    if extras and rank == 0:
        rec["loss_hbm"] = _loss_hbm(dev, peaks)
        rec["c1_batch4"] = {"256": _small_batch(a, dev, 4, 256), "128": _small_batch(a, dev, 4, 128)}
        if world == 1:
            # the same step at 128 and 256 images: the ~10 us of fixed cost per conv launch amortise with the batch
            fpi = rec.get("roofline", {}).get("flops_per_step", 0.0) / max(1, a.batch)
            sweep = {}
            for b, st in ((128, 20), (256, 12)):
                r = _small_batch(a, dev, b, a.size, steps=st)
                if fpi and peaks.get("tensor"):
                    r["whole_step_frac_of_peak"] = r["images_per_sec"] * fpi / (peaks["tensor"] * 1e12)
                sweep[str(b)] = r
            rec["batch_sweep"] = sweep
    if extras:
        if world > 1:
            dist.barrier()
        c5 = _dp_c5(a, dev, world, rank, peaks)
        if rank == 0:
            rec["dp_c5"] = c5
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        rec["cpu_baseline"] = cpu_baseline(a, a.cpu_seconds)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        _emit(rec)


# ------------------------------------------------ CPU baseline / reference arm ----
def _reference_stepper(a):
    """(step(images, masks), kind, what): one training step of the reference path on the host.

    kind "reference": the reference's OWN modules staged under oracle/_ref by oracle/make_ref.py --
    advanced_models.UNetWithBackbone, losses.get_loss_function, torch.optim.Adam as
    cross_validation.py:93-97 builds it, and train.train_epoch (train.py:17-68: forward, loss,
    backward, step, calculate_metrics) over a one-batch loader.  kind "port": oracle/train_step.py,
    the functional restatement on the same ATen CPU kernels, when oracle/_ref is absent."""
    import torch
    from oracle import ref_loader
    if ref_loader.available() and "advanced_models" not in sys.modules:
        ref = ref_loader.reference_modules()
        torch.manual_seed(1)
        model = ref.advanced_models.UNetWithBackbone(n_classes=1, backbone=a.backbone, pretrained=False,
                                                     use_attention=not a.no_attention)
        crit = ref.losses.get_loss_function({"loss_fn": a.loss})
        opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=1e-5)
        cpu = torch.device("cpu")

        def step(images, masks):
            return ref.train.train_epoch(model, [(images, masks)], opt, crit, cpu)

        return step, "reference", "the reference's own train.train_epoch / UNetWithBackbone / losses (oracle/_ref)"
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
    trainer = OracleTrainer(sd, a.backbone, not a.no_attention, a.loss)
    return trainer.step, "port", "oracle/train_step.py (functional port: forward, loss, backward, Adam, metrics)"


def _synth_cpu(batch, size, seed):
    """b2seg.synth.cell_tiles without importing the package's GPU modules (pure torch CPU code)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_b2seg_synth", os.path.join(PKG, "b2seg", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.cell_tiles(batch, size, seed=seed)


def _cpu_steps(a, batch, steps, warmup, budget_s=None):
    """Times `steps` reference steps on batches of `batch` images; returns (seconds, steps done, ...)."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    step, kind, what = _reference_stepper(a)
    imgs, masks = _synth_cpu(batch, a.size, 1234)
    for _ in range(warmup):
        step(imgs, masks)
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        step(imgs, masks)
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s and done >= 2:
            break
    return time.perf_counter() - t0, done, cores, kind, what


def cpu_baseline(a, budget_s):
    """The reference arm on a bounded sample, in its OWN process (the reference's modules carry the
    same names as the B200 ones: they cannot share an interpreter)."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "1000000", "--warmup", "1",
           "--batch", str(a.batch), "--size", str(a.size), "--backbone", a.backbone, "--loss", a.loss,
           "--cpu-seconds", str(budget_s)] + (["--no-attention"] if a.no_attention else [])
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=budget_s * 4 + 240, env=env)
        line = [ln for ln in out.stdout.splitlines() if ln.startswith("{")][-1]
        rec = json.loads(line)
        cb = rec["cpu_baseline"]
        cb["ms_per_step"] = rec["ms_per_step"]
        return cb
    except Exception as e:  # the bench line must still come out
        return {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "unavailable",
                "sample": f"cpu baseline failed: {type(e).__name__}: {e}"}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return  # one CPU arm per box: the host cores are shared
    batch = min(a.batch, 8)
    steps = max(1, a.steps)
    budget = a.cpu_seconds if a.steps >= 1000000 else 240.0
    sec, done, cores, kind, what = _cpu_steps(a, batch, steps=steps, warmup=max(0, a.warmup), budget_s=budget)
    v = batch * done / sec
    sample = (f"each step = {batch} images ({a.size}x{a.size}), a bounded sample of the "
              f"{a.batch}-image batch; {done} timed steps after {a.warmup} warm-up; fp32 torch CPU, {cores} threads; {what}")
    rec = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": max(world, a.gpus),
        "steps": done, "warmup": a.warmup, "ms_per_step": sec / done * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(workload(a), global_batch=a.batch * max(world, a.gpus), parallelism="cpu"),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
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
