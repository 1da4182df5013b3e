"""Instance sharding across the GPUs of one box.

MPC instances are independent, so a sweep is split into contiguous slices, one per rank (one process
per GPU, one engine handle / workspace / stream each).  There is NO collective on the solve path; the
only communication is the optional final gather of the small per-instance results
(``cost / iters / ok / f_pred`` and, when asked, trajectories) onto rank 0, through
``torch.distributed`` (NCCL over NVLink on the GPU box, gloo in CPU tests).

The reference has no batched mode (one instance per 5 ms tick, ``run_classical.py:431``); the scenario
sweep it runs sequentially (``run_classical.py:30-39`` seeds) is what gets sharded here.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional, Sequence, Tuple

import numpy as np

from .batch import BatchInputs


def shard_bounds(total: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous slice [lo, hi) of `total` instances owned by `rank`; sizes differ by at most one."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} / world {world}")
    if total < 0:
        raise ValueError("total must be >= 0")
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_sizes(total: int, world: int) -> np.ndarray:
    return np.array([shard_bounds(total, world, r)[1] - shard_bounds(total, world, r)[0] for r in range(world)], dtype=np.int64)


def scenario_of_instance(total: int, scenarios: Sequence[str]) -> np.ndarray:
    """Scenario index of every instance of a sweep laid out scenario-major (config 4: 4 x 16384)."""
    per = shard_sizes(total, len(scenarios))
    return np.repeat(np.arange(len(scenarios)), per)


def shard_inputs(inp: BatchInputs, world: int, rank: int) -> BatchInputs:
    lo, hi = shard_bounds(inp.B, world, rank)
    return inp.slice(np.arange(lo, hi))


def gather_results(local: Dict[str, np.ndarray], total: int, world: int, rank: int, names: Optional[Sequence[str]] = None,
                   dist=None, device=None) -> Optional[Dict[str, np.ndarray]]:
    """Gather per-instance results of every rank onto rank 0 in instance order.

    `local[name]` has leading dimension = this rank's shard size.  Uses all_gather on tensors padded to the
    largest shard (works for both gloo and nccl; no data-path collective ever precedes this call).
    Returns the full arrays on rank 0 and None elsewhere.  `dist=None` (single process) returns `local`.
    """
    names = list(names if names is not None else local.keys())
    if world == 1 or dist is None:
        return {n: np.asarray(local[n]) for n in names}
    import torch

    sizes = shard_sizes(total, world)
    cap = int(sizes.max())
    out = {} if rank == 0 else None
    for n in names:
        a = np.ascontiguousarray(local[n])
        if a.shape[0] != sizes[rank]:
            raise ValueError(f"{n}: leading dimension {a.shape[0]} != shard size {sizes[rank]}")
        pad = np.zeros((cap,) + a.shape[1:], dtype=a.dtype)
        pad[: a.shape[0]] = a
        t = torch.from_numpy(pad)
        if device is not None:
            t = t.to(device)
        bufs = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(bufs, t)
        if rank == 0:
            out[n] = np.concatenate([bufs[r][: int(sizes[r])].cpu().numpy() for r in range(world)], axis=0)
    return out


def solve_sharded(inp: BatchInputs, world: int, rank: int, solve: Callable[[BatchInputs], Dict[str, np.ndarray]],
                  names: Sequence[str] = ("cost", "iters", "ok", "f_pred"), dist=None, device=None):
    """Slice -> local solve (`solve` = e.g. ``lambda s: engine.solve_host(s, iters)``) -> gather on rank 0."""
    shard = shard_inputs(inp, world, rank)
    local = solve(shard) if shard.B > 0 else {}
    if shard.B == 0:
        raise ValueError("empty shard: fewer instances than ranks")
    return gather_results(local, inp.B, world, rank, names, dist=dist, device=device)
