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


def gather_results(local: Dict[str, object], total: int, world: int, rank: int, names: Optional[Sequence[str]] = None,
                   dist=None, device=None, to_numpy: bool = True):
    """Gather per-instance results of every rank in instance order (north_star: "only a final result gather over NVLink").

    `local[name]`: numpy array or torch tensor whose leading dimension is this rank's shard size.  Torch tensors stay
    where they are: a CUDA tensor is padded to the largest shard ON ITS DEVICE and all-gathered by NCCL from device memory
    (no host round trip); numpy arrays go through `device` (None: CPU tensors, the gloo test path).  No data-path
    collective ever precedes this call.  Rank 0 gets the full arrays (numpy when `to_numpy`, else torch tensors on the
    gathering device), the other ranks None.  `dist=None` (single process) returns `local`."""
    names = list(names if names is not None else local.keys())
    if world == 1 or dist is None:
        return {n: (np.asarray(local[n].cpu() if hasattr(local[n], "cpu") else local[n]) if to_numpy else local[n]) for n in names}
    import torch

    sizes = shard_sizes(total, world)
    cap = int(sizes.max())
    # ONE collective: every array's rows are packed side by side into one byte matrix [cap, row_bytes] (the per-instance
    # results are tens of bytes; four separate all-gathers cost four collective latencies for half a megabyte)
    parts, meta = [], []
    for n in names:
        a = local[n]
        t = a if torch.is_tensor(a) else torch.from_numpy(np.ascontiguousarray(a))
        if device is not None and t.device != torch.device(device):
            t = t.to(device)
        if t.shape[0] != sizes[rank]:
            raise ValueError(f"{n}: leading dimension {t.shape[0]} != shard size {sizes[rank]}")
        rows = t.shape[0]
        b = t.contiguous().reshape(rows, -1).view(torch.uint8)   # [rows, bytes per instance of this array]
        parts.append(b)
        meta.append((n, t.dtype, tuple(t.shape[1:]), b.shape[1]))
    packed = torch.cat(parts, dim=1) if len(parts) > 1 else parts[0]
    if packed.shape[0] != cap:
        pad = torch.zeros((cap, packed.shape[1]), dtype=torch.uint8, device=packed.device)
        pad[: packed.shape[0]] = packed
        packed = pad
    packed = packed.contiguous()
    full = torch.empty((world * cap, packed.shape[1]), dtype=torch.uint8, device=packed.device)
    dist.all_gather_into_tensor(full, packed)
    if rank != 0:
        return None
    keep = torch.cat([full[r * cap: r * cap + int(sizes[r])] for r in range(world)], dim=0) if int(sizes.min()) != cap else full
    out, col = {}, 0
    for n, dtype, tail, nb in meta:
        v = keep[:, col: col + nb].contiguous().view(dtype).reshape((keep.shape[0],) + tail)
        col += nb
        out[n] = v.cpu().numpy() if to_numpy else v
    return out


def solve_sharded(inp: BatchInputs, world: int, rank: int, solve: Callable[[BatchInputs], Dict[str, np.ndarray]],
                  names: Sequence[str] = ("cost", "iters", "ok", "f_pred"), dist=None, device=None):
    """Slice -> local solve (`solve` = e.g. ``lambda s: engine.solve_host(s, iters)``) -> gather on rank 0.
    Every rank checks the WHOLE partition before anybody solves or enters the collective: a rank raising alone on its
    empty shard would leave the others waiting in all_gather."""
    if inp.B < world:
        raise ValueError(f"{inp.B} instances over {world} ranks: some rank would get an empty shard")
    shard = shard_inputs(inp, world, rank)
    return gather_results(solve(shard), inp.B, world, rank, names, dist=dist, device=device)
