"""How fast can N ranks write results into this host's memory at once?  Each rank copies a device buffer to pinned host
memory in a loop (plain cudaMemcpyAsync, as the engine's host API does); rank 0 prints the per-rank and aggregate rates.
    python -m torch.distributed.run --nproc-per-node N tools/d2h_probe.py [--gb 2] [--reps 8] [--chunks 1]"""
import argparse
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=2.0)
    ap.add_argument("--reps", type=int, default=8)
    ap.add_argument("--chunks", type=int, default=1, help="concurrent copies per rank, each on its own stream")
    a = ap.parse_args()
    rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    n = int(a.gb * 1e9 / 8 / a.chunks)
    dev = [torch.zeros(n, dtype=torch.float64, device="cuda") for _ in range(a.chunks)]
    host = [torch.empty(n, dtype=torch.float64).pin_memory() for _ in range(a.chunks)]
    streams = [torch.cuda.Stream() for _ in range(a.chunks)]
    for direction in ("d2h", "h2d"):
        for c in range(a.chunks):
            with torch.cuda.stream(streams[c]):
                (host[c].copy_(dev[c], non_blocking=True) if direction == "d2h" else dev[c].copy_(host[c], non_blocking=True))
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(a.reps):
            for c in range(a.chunks):
                with torch.cuda.stream(streams[c]):
                    (host[c].copy_(dev[c], non_blocking=True) if direction == "d2h" else dev[c].copy_(host[c], non_blocking=True))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        rate = torch.tensor([n * 8 * a.chunks * a.reps / dt * 1e-9], device="cuda", dtype=torch.float64)
        rates = [torch.zeros_like(rate) for _ in range(world)]
        if world > 1:
            dist.all_gather(rates, rate)
        else:
            rates = [rate]
        if rank == 0:
            r = [float(x.item()) for x in rates]
            print(f"{direction} world={world} chunks={a.chunks} {a.gb} GB x {a.reps}: per rank {min(r):.1f} .. {max(r):.1f} GB/s, aggregate {sum(r):.1f} GB/s", flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
