"""CPU ranks: the shared-memory all-reduce (csrc/runtime/shm_comm.cc) against torch.distributed /
gloo on the same box, fp32 sums from 1 KiB to 64 MiB.  Times are the max over ranks of the mean
per call; busBW uses the NCCL-tests convention (algBW x 2(N-1)/N).

    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 bench/shm_allreduce.py
"""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200.parallel import shm  # noqa: E402


def timed(fn, iters, group_barrier):
    for _ in range(3):
        fn()
    group_barrier()
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    dt = (time.perf_counter() - t0) / iters
    t = torch.tensor([dt])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--max-mb", type=int, default=64)
    args = ap.parse_args()
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    arena = shm.ShmArena((args.max_mb << 20) + 4096)
    rows = []
    nbytes = 1024
    while nbytes <= args.max_mb << 20:
        n = nbytes // 4
        t = arena.tensor(n, torch.float32) if nbytes == 1024 else arena_view(arena, n)
        u = torch.zeros(n)
        iters = max(5, min(300, (32 << 20) // nbytes))
        row = {"bytes": nbytes}
        for name, fn in (("shm_oneshot", lambda: arena.all_reduce(t, algo=shm.ALGO_ONESHOT)),
                         ("shm_twoshot", lambda: arena.all_reduce(t, algo=shm.ALGO_TWOSHOT)),
                         ("gloo", lambda: dist.all_reduce(u))):
            dt = timed(fn, iters, dist.barrier)
            row[name + "_us"] = dt * 1e6
            row[name + "_busbw_gbs"] = nbytes / dt * 2 * (world - 1) / world / 1e9
        rows.append(row)
        if rank == 0:
            print("{:>10} B  one-shot {:9.1f} us  two-shot {:9.1f} us  gloo {:9.1f} us".format(
                nbytes, row["shm_oneshot_us"], row["shm_twoshot_us"], row["gloo_us"]), flush=True)
        nbytes *= 4
    if rank == 0 and args.out:
        with open(args.out, "w") as f:
            json.dump({"world": world, "cpus": os.cpu_count(), "rows": rows}, f, indent=1)


_big = {}


def arena_view(arena, n):
    """One big tensor handed out once; smaller sizes are slices of it."""
    if "t" not in _big:
        _big["t"] = arena.tensor((arena.free_bytes - 128) // 4, torch.float32)
    return _big["t"][:n]


if __name__ == "__main__":
    main()
