"""Multi-GPU host logic on CPU: unit sharding (folds / grid configs, no data-path collective),
the sharded CrossValidator's merged result structure, and the data-parallel gradient hook --
world_size 2 over gloo, mirroring what runs one-rank-per-GPU over NCCL on the box."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG = os.path.join(ROOT, "image-segmentation-project_b200")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn_name, out_dir):
    for p in (PKG, ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        res = globals()[fn_name](rank, world)
        torch.save(res, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _spawn(fn_name, tmp_path, world=2):
    mp.spawn(_worker, args=(world, _free_port(), fn_name, str(tmp_path)), nprocs=world, join=True)
    return [torch.load(os.path.join(str(tmp_path), f"r{r}.pt"), weights_only=False) for r in range(world)]


# ------------------------------------------------------------------ assign ----
def test_assign_partitions_every_unit_once():
    from b2seg import shard
    for n in (0, 1, 5, 20, 48):
        for w in (1, 2, 4, 8):
            for costs in (None, [((7 * i) % 5) + 1.0 for i in range(n)]):
                parts = [shard.assign(n, r, w, costs) for r in range(w)]
                flat = sorted(i for p in parts for i in p)
                assert flat == list(range(n))
    assert shard.assign(5, 0, 2) == [0, 2, 4]
    with pytest.raises(ValueError):
        shard.assign(4, 2, 2)
    with pytest.raises(ValueError):
        shard.assign(4, 0, 2, costs=[1.0])


def test_assign_balances_mixed_costs():
    from b2seg import shard
    # the model_configuration grid mixes resnet34 (1x) and resnet50 (4.5x) units
    costs = [4.5] * 8 + [1.0] * 32
    loads = [sum(costs[i] for i in shard.assign(len(costs), r, 8, costs)) for r in range(8)]
    assert max(loads) - min(loads) <= 1.0
    rr = [sum(costs[i] for i in shard.assign(len(costs), r, 8)) for r in range(8)]
    assert max(loads) <= max(rr)


def test_run_units_sequential_matches_loop():
    from b2seg import shard
    units = list(range(7))
    assert shard.run_units(units, lambda u: {"v": u * u}) == [{"v": u * u} for u in units]
    with pytest.raises(ZeroDivisionError):
        shard.run_units([1, 0], lambda u: 1 // u)


# --------------------------------------------------------- gloo, world 2 ----
def _job_run_units(rank, world):
    from b2seg import shard
    ran = []

    def fn(u):
        ran.append(u)
        return {"unit": u, "rank": rank, "iou": 0.5 + 0.01 * u}

    plain = shard.run_units(list(range(9)), fn)
    ran_plain = list(ran)
    del ran[:]
    costed = shard.run_units(list(range(9)), fn, costs=[9, 1, 1, 1, 1, 1, 1, 1, 1])
    err = None
    try:
        shard.run_units([1, 2, 0, 4], lambda u: 8 // u)
    except RuntimeError as e:
        err = str(e)
    return {"plain": plain, "ran_plain": ran_plain, "costed": costed, "ran_costed": list(ran), "err": err}


def test_run_units_gloo_world2(tmp_path):
    r0, r1 = _spawn("_job_run_units", tmp_path)
    # every rank ends with all results in unit order; each unit ran on exactly one rank
    assert r0["plain"] == r1["plain"]
    assert [d["unit"] for d in r0["plain"]] == list(range(9))
    assert [d["rank"] for d in r0["plain"]] == [u % 2 for u in range(9)]
    assert sorted(r0["ran_plain"] + r1["ran_plain"]) == list(range(9))
    assert r0["costed"] == r1["costed"]
    assert sorted(r0["ran_costed"] + r1["ran_costed"]) == list(range(9))
    assert r0["ran_costed"] == [0]  # LPT: the 9-cost unit alone on rank 0 (ranks 1 takes 8 x 1)
    # a failing unit surfaces on BOTH ranks instead of hanging the gather
    assert r0["err"] and r1["err"] and "ZeroDivisionError" in r0["err"] and r0["err"] == r1["err"]


def _fake_cv(cv_mod, n_images=23):
    """CrossValidator over fake paths whose 'training' is a pure function of the fold."""
    cv_mod.load_original_data = lambda data_dir, image_type: {
        "image_paths": [f"img_{i:03d}.tif" for i in range(n_images)],
        "mask_paths": [f"msk_{i:03d}.tif" for i in range(n_images)]}
    cv = cv_mod.CrossValidator("nowhere", "W", n_splits=5, random_state=42, augmentations_per_image=0,
                               verbose=False)

    def fake_train(model_class, config, tr_i, tr_m, va_i, va_m):
        key = sum(int(p[4:7]) for p in va_i) + 1000 * config["learning_rate"]
        return {"precision": key % 7 / 7, "recall": key % 5 / 5, "f1": key % 3 / 3,
                "iou": (key % 11) / 11, "accuracy": 0.9, "loss": 1.0 / (1 + key)}

    cv.train_single_model = fake_train
    return cv


CONFIGS = [(None, {"name": f"cfg{i}", "learning_rate": lr, "backbone": bb, "num_epochs": 2})
           for i, (lr, bb) in enumerate([(1e-3, "resnet34"), (1e-4, "resnet50"), (5e-4, "resnet34")])]


def _job_cv(rank, world):
    import cross_validation as cv_mod
    cv = _fake_cv(cv_mod)
    out = cv.compare_multiple_models(CONFIGS)
    single = cv.cross_validate_single_model(None, CONFIGS[0][1], indices=list(range(3, 20)))
    for res in [single] + list(out["individual_results"].values()):
        for fr in res["fold_results"]:
            fr.pop("training_time")
    return {"cmp": out, "single": single}


def test_sharded_cross_validation_equals_sequential(tmp_path):
    for p in (PKG, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    seq = _job_cv(0, 1)  # no process group in this process: the reference's sequential loop
    r0, r1 = _spawn("_job_cv", tmp_path)
    for got in (r0, r1):
        assert got["cmp"]["cv_folds_used"] == 5
        assert list(got["cmp"]["individual_results"]) == ["cfg0", "cfg1", "cfg2"]
        assert got["cmp"]["individual_results"] == seq["cmp"]["individual_results"]
        assert got["cmp"]["comparison_summary"] == seq["cmp"]["comparison_summary"]
        assert got["single"] == seq["single"]
    s = seq["cmp"]["comparison_summary"]["cfg0"]
    assert set(s) == {f"{m}_{k}" for m in ("precision", "recall", "f1", "iou", "accuracy", "loss")
                      for k in ("mean", "std")}


class _Toy(torch.nn.Module):
    def __init__(self, seed):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.w = torch.nn.Parameter(torch.randn(5, 3, generator=g))
        self.register_buffer("running", torch.randn(4, generator=g))
        self._grad_hook = None


def _job_dp(rank, world):
    from b2seg import dataparallel
    model = _Toy(seed=10 + rank)  # replicas start different on purpose
    dataparallel.attach(model, world)
    gflat = torch.arange(8, dtype=torch.float32) * (rank + 1)
    model._grad_hook(gflat)
    # the bucketed form: slices go out one by one (async), finish() waits for all and averages
    g2 = torch.arange(8, dtype=torch.float32) * (rank + 1)
    for lo, hi in ((5, 8), (2, 5), (0, 2)):
        model._grad_hook.bucket(g2[lo:hi])
    model._grad_hook.finish(g2)
    dataparallel.detach(model)
    return {"w": model.w.detach().clone(), "running": model.running.clone(), "g": gflat, "g2": g2,
            "hook_cleared": model._grad_hook is None}


def test_dataparallel_hook_gloo_world2(tmp_path):
    r0, r1 = _spawn("_job_dp", tmp_path)
    ref = _Toy(seed=10)
    assert torch.equal(r0["w"], ref.w.detach()) and torch.equal(r1["w"], ref.w.detach())
    assert torch.equal(r0["running"], ref.running) and torch.equal(r1["running"], ref.running)
    want = torch.arange(8, dtype=torch.float32) * 1.5  # mean of (1x, 2x)
    assert torch.equal(r0["g"], want) and torch.equal(r1["g"], want)
    assert torch.equal(r0["g2"], want) and torch.equal(r1["g2"], want)
    assert r0["hook_cleared"] and r1["hook_cleared"]
