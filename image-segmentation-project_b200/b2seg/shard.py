"""Shard independent training units over the GPUs of one box (north_star: "cross-validation
folds and grid-search configs shard one per GPU with no communication").

A *unit* is one (config, fold) training run -- the body of the reference's nested loops
(cross_validation.py:220 x 229, model_configuration_experiment.py:51,
forward_selection_integration.py:509 x 437).  Units never exchange tensors: the only
communication is one host-side gather of the small python result dicts at the end, after which
every rank holds the results of ALL units in their sequential order, so callers can build exactly
the structure the reference builds (cross_validation.py:294-298).

One process per GPU (torchrun / torch.distributed).  Without an initialised process group every
function degenerates to the sequential loop of the reference.
"""
from __future__ import annotations

import datetime
import os
from typing import Callable, List, Sequence

import torch
import torch.distributed as dist


def world() -> tuple:
    """(rank, world_size) of the current process; (0, 1) when not launched distributed."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


# Units of a grid can differ by hours (resnet50 vs resnet34, early stopping): the result gather
# must not sit under NCCL's 10-minute watchdog.  It runs over a gloo (CPU) group with this timeout;
# NCCL is kept for tensors only (the data-parallel gradient all-reduce).
RESULT_TIMEOUT = datetime.timedelta(hours=float(os.environ.get("B2SEG_SHARD_TIMEOUT_HOURS", "48")))
_result_group = None


def init_from_env(backend: str | None = None) -> tuple:
    """Join the torchrun rendezvous if RANK/WORLD_SIZE are set (idempotent); bind this process
    to its GPU (LOCAL_RANK).  Returns (rank, world_size).  CrossValidator / ModelComparator call
    this, so `torchrun --nproc-per-node N <experiment script>.py` shards without any change to
    the script."""
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if ws > 1 and "RANK" in os.environ:
        local = int(os.environ.get("LOCAL_RANK", "0"))
        if torch.cuda.is_available():
            torch.cuda.set_device(local % torch.cuda.device_count())
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            if backend == "nccl":
                dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()),
                                        timeout=RESULT_TIMEOUT)
            else:
                dist.init_process_group(backend, timeout=RESULT_TIMEOUT)
    return world()


def result_group():
    """The process group the small python result dicts travel over: the default group when it is
    gloo already, otherwise a gloo group created next to the NCCL one (collective: every rank
    reaches this from run_units at the same point)."""
    global _result_group
    if not dist.is_initialized():
        return None
    if dist.get_backend() == "gloo":
        return None  # default group
    if _result_group is None:
        _result_group = dist.new_group(backend="gloo", timeout=RESULT_TIMEOUT)
    return _result_group


def assign(n_units: int, rank: int, world_size: int, costs: Sequence[float] | None = None) -> List[int]:
    """Indices of the units `rank` runs.  Round-robin by default (unit i -> rank i % world);
    with `costs` (e.g. epochs x images), longest-processing-time-first onto the least loaded rank
    so a grid mixing resnet34 and resnet50 configs finishes together.  Deterministic, identical
    on every rank, every unit assigned exactly once."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world {rank}/{world_size}")
    if costs is None:
        return list(range(rank, n_units, world_size))
    if len(costs) != n_units:
        raise ValueError("costs must have one entry per unit")
    load = [0.0] * world_size
    mine = []
    for i in sorted(range(n_units), key=lambda j: (-float(costs[j]), j)):
        r = min(range(world_size), key=lambda q: (load[q], q))
        load[r] += float(costs[i])
        if r == rank:
            mine.append(i)
    return sorted(mine)


def run_units(units: Sequence, fn: Callable, costs: Sequence[float] | None = None, group=None) -> list:
    """results[i] = fn(units[i]) for every unit, each computed on exactly one rank, returned in
    unit order on EVERY rank.  `fn` must return something picklable and small (metric dicts).
    An exception in any unit is re-raised on all ranks after the gather (no rank hangs)."""
    rank, ws = world()
    mine = assign(len(units), rank, ws, costs)
    local, err = {}, None
    for i in mine:
        try:
            local[i] = fn(units[i])
        except Exception as e:  # keep the collective below matched on every rank
            err = (i, f"{type(e).__name__}: {e}")
            if ws == 1:
                raise
            break
    if ws == 1:
        return [local[i] for i in range(len(units))]
    gathered = [None] * ws
    dist.all_gather_object(gathered, (local, err), group=group if group is not None else result_group())
    merged = {}
    for r, (part, e) in enumerate(gathered):
        if e is not None:
            raise RuntimeError(f"unit {e[0]} failed on rank {r}: {e[1]}")
        merged.update(part)
    missing = [i for i in range(len(units)) if i not in merged]
    if missing:
        raise RuntimeError(f"units {missing} produced no result")
    return [merged[i] for i in range(len(units))]
