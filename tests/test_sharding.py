"""Multi-GPU host logic on CPU: slice bounds, scenario layout, and the world_size-2 gather over gloo
(instances shard with no data-path collective; only the result gather communicates).
The per-shard "solve" here is the oracle (test infrastructure) standing in for the engine: what is under
test is that sharding + gather reproduces the single-process result, instance for instance."""
from __future__ import annotations

import os
import socket
import sys

import numpy as np
import pytest

from ffmpc_cases import make_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_bounds_cover_and_balance():
    from franka_force_feedback_mpc_b200.sharding import scenario_of_instance, shard_bounds, shard_sizes

    for total in (0, 1, 7, 16384, 65536, 65537):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_bounds(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sz = shard_sizes(total, world)
            assert sz.sum() == total and sz.max() - sz.min() <= 1
    with pytest.raises(ValueError):
        shard_bounds(10, 2, 2)
    sc = scenario_of_instance(65536, ("tilted_5", "tilted_10", "tilted_15", "actuation_uncertainty"))
    assert sc.size == 65536 and np.bincount(sc).tolist() == [16384] * 4
    # config 4 on 8 GPUs: every rank's slice holds exactly one scenario
    for r in range(8):
        lo, hi = shard_bounds(65536, 8, r)
        assert np.unique(sc[lo:hi]).size == 1


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist

    from franka_force_feedback_mpc_b200.sharding import solve_sharded
    from oracle import oracle as orc

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        cfg, am, inp = make_case(True, 1, 5, 6, 77, 0.6)   # 5 instances over 2 ranks: ragged shards (3 + 2)
        names = ("cost", "iters", "ok", "f_pred", "us")
        res = solve_sharded(inp, world, rank, lambda s: orc.solve_batch(cfg, am, s, 3, True, nthreads=1), names, dist=dist)
        if rank == 0:
            q.put({k: v for k, v in res.items()})
        else:
            assert res is None
        # torch tensors in (the device-resident form of the gather; CPU tensors under gloo), torch tensors out
        import torch

        from franka_force_feedback_mpc_b200.sharding import gather_results, shard_inputs
        loc = orc.solve_batch(cfg, am, shard_inputs(inp, world, rank), 3, True, nthreads=1)
        tt = gather_results({k: torch.from_numpy(loc[k]) for k in ("cost", "iters")}, inp.B, world, rank, dist=dist, to_numpy=False)
        if rank == 0:
            assert torch.is_tensor(tt["cost"]) and np.array_equal(tt["cost"].numpy(), res["cost"]) and np.array_equal(tt["iters"].numpy(), res["iters"])
        # fewer instances than ranks: EVERY rank raises before any collective (no deadlock)
        try:
            solve_sharded(inp.slice(np.arange(1)), world, rank, lambda s: {}, names, dist=dist)
            raise AssertionError("expected ValueError")
        except ValueError:
            pass
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_world2_gloo_gather_matches_single_process(orc):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=240)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    cfg, am, inp = make_case(True, 1, 5, 6, 77, 0.6)
    ref = orc.solve_batch(cfg, am, inp, 3, True, nthreads=1)
    for k in ("cost", "iters", "ok", "f_pred", "us"):
        assert res[k].shape == ref[k].shape
        np.testing.assert_array_equal(res[k], ref[k], err_msg=k)   # same code, same inputs: bit-identical
