"""Gradient all-reduce sweep (BASELINE.json config 5): bus bandwidth of the hand-written
peer-memory all-reduce (one-shot / two-shot / NVLS) against NCCL, 1 KB - 1 GB, fp32 sum, plus a
correctness check of every schedule (plain reduce, fused Adam) against torch.distributed.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
        bench/allreduce_sweep.py [--check-only] [--max-mb 1024]

busBW = algBW * 2(N-1)/N (NCCL-tests convention); times are CUDA-event device times, max over
ranks, L2 flushed between iterations by rotating through buffers larger than L2 where possible.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200 import ops  # noqa: E402
from cloud_b200.parallel import comm as C  # noqa: E402
from cloud_b200.parallel.comm import Communicator  # noqa: E402
from cloud_b200.ops.optim import _reference_update, make_hyper, optim_tick  # noqa: E402


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


def check(comm, dev, world, rank, results):
    """Every schedule x dtype: reduce-only and fused Adam vs NCCL + reference update."""
    n = 8 * world * 4099          # odd-ish size, multiple of 8*world
    ok_all = True
    for dtype in (torch.float32, torch.bfloat16):
        item = 4 if dtype == torch.float32 else 2
        gbuf, pbuf = comm.alloc_pair(n * item, n * item)
        ctx = comm.make_ctx(gbuf, pbuf)
        g_arena, p_arena = gbuf.view(dtype, n), pbuf.view(dtype, n)
        algos = [("oneshot", C.ALGO_ONESHOT), ("twoshot", C.ALGO_TWOSHOT)]
        if gbuf.mc_ptr and pbuf.mc_ptr:
            algos.append(("nvls", C.ALGO_NVLS))
        for name, algo in (("ll1", C.ALGO_LL1), ("ll2", C.ALGO_LL2)):
            if n * item <= comm.ll_capacity(algo):
                algos.append((name, algo))
        hyper = make_hyper(dev, lr=1e-2, grad_scale=1.0 / world)
        for name, algo in algos:
            # ---- plain reduction -----------------------------------------------------------
            torch.manual_seed(100 + rank)
            g = torch.randn(n, device=dev).to(dtype)
            g_arena.copy_(g)
            want = g.float().clone()
            dist.all_reduce(want)
            want = want / world
            torch.cuda.synchronize()
            dist.barrier()
            comm.fused(ctx, algo=algo, opt=0, grad_dtype=dtype, param_dtype=dtype, off=0, length=n,
                       master=None, s1=None, s2=None, state_off=0, hyper=hyper, zero_grads=True)
            torch.cuda.synchronize()
            dist.barrier()
            got = p_arena.float()
            tol = 1e-5 if dtype == torch.float32 else 2e-2
            err = (got - want).abs().max().item()
            zeroed = g_arena.float().abs().max().item() == 0.0
            ok = err <= tol * max(1.0, want.abs().max().item()) and zeroed
            # replicas must agree bit for bit
            ref0 = got.clone()
            dist.broadcast(ref0, src=0)
            same = torch.equal(ref0, got)
            results.append({"check": f"reduce/{name}/{str(dtype)[6:]}", "max_err": err,
                            "grads_zeroed": zeroed, "replicas_identical": same, "ok": ok and same})
            ok_all &= ok and same
            # ---- fused Adam, 3 steps ---------------------------------------------------------
            sharded = algo in C.SHARDED_ALGOS
            n_state = n // world if sharded else n
            master = torch.zeros(n_state, device=dev)
            m = torch.zeros(n_state, device=dev)
            v = torch.zeros(n_state, device=dev)
            torch.manual_seed(7)
            p0 = torch.randn(n, device=dev)
            p_arena.copy_(p0.to(dtype))
            lo = rank * n_state if sharded else 0
            master.copy_(p_arena[lo:lo + n_state].float())
            r_p, r_m, r_v = p_arena.float().cpu().clone(), torch.zeros(n), torch.zeros(n)
            h2 = make_hyper(dev, lr=1e-2, grad_scale=1.0 / world)
            for step in range(3):
                torch.manual_seed(1000 * step + rank)
                g = torch.randn(n, device=dev).to(dtype)
                g_arena.copy_(g)
                gs = g.float().clone()
                dist.all_reduce(gs)
                optim_tick(h2)
                torch.cuda.synchronize()
                dist.barrier()
                comm.fused(ctx, algo=algo, opt=ops.OPT_CODES["adam"], grad_dtype=dtype,
                           param_dtype=dtype, off=0, length=n, master=master, s1=m, s2=v,
                           state_off=0, hyper=h2, zero_grads=True)
                torch.cuda.synchronize()
                dist.barrier()
                r_p = _reference_update(1, r_p, gs.cpu(), r_m, r_v, h2.cpu())
            got = p_arena.float().cpu()
            tol = 1e-4 if dtype == torch.float32 else 2e-2
            err = (got - r_p).abs().max().item()
            ok = err <= tol * max(1.0, r_p.abs().max().item())
            results.append({"check": f"adam/{name}/{str(dtype)[6:]}", "max_err": err, "ok": ok})
            ok_all &= ok
        # ---- broadcast ----------------------------------------------------------------------
        p_arena.fill_(float(rank + 1))
        torch.cuda.synchronize()
        dist.barrier()
        comm.broadcast_arena(ctx, dtype, 0, n, root=0)
        torch.cuda.synchronize()
        dist.barrier()
        ok = bool((p_arena.float() == 1.0).all().item())
        results.append({"check": f"broadcast/{str(dtype)[6:]}", "ok": ok})
        ok_all &= ok
    return ok_all


def stress(comm, dev, world, rank, results, iters=48):
    """Skewed arrival: back-to-back fused all-reduces with NO host barrier
    in between, every rank delaying both its host thread and its stream by a different random
    amount before each call, schedules and lengths alternating.  Exercises the per-(block, source)
    epoch flags and the closing barrier of the kernels when a fast rank is already two calls ahead
    of a slow one."""
    import random
    import time
    n_max = 8 * world * 8209
    gbuf, pbuf = comm.alloc_pair(n_max * 4, n_max * 4)
    ctx = comm.make_ctx(gbuf, pbuf)
    g_arena, p_arena = gbuf.view(torch.float32, n_max), pbuf.view(torch.float32, n_max)
    algos = [C.ALGO_ONESHOT, C.ALGO_TWOSHOT] + ([C.ALGO_NVLS] if gbuf.mc_ptr and pbuf.mc_ptr else [])
    if comm.ll_capacity(C.ALGO_LL1) >= n_max * 4:
        algos += [C.ALGO_LL1, C.ALGO_LL2, C.ALGO_LL1, C.ALGO_LL2]     # back to back: exercises the parities
    hyper = make_hyper(dev, lr=1e-2, grad_scale=1.0 / world)
    rng = random.Random(1234 + 17 * rank)            # different delays on every rank
    lens = [8 * world * k for k in (1, 33, 1027, 8209)]
    bad = 0
    dist.barrier()
    for it in range(iters):
        n = lens[it % len(lens)]
        algo = algos[it % len(algos)]
        g_arena[:n].fill_(float((rank + 1) * (it + 1)))
        time.sleep(rng.random() * 0.003)
        torch.cuda._sleep(int(rng.random() * 3e6))    # up to ~1.5 ms of stream delay
        comm.fused(ctx, algo=algo, opt=0, grad_dtype=torch.float32, param_dtype=torch.float32, off=0,
                   length=n, master=None, s1=None, s2=None, state_off=0, hyper=hyper, zero_grads=True)
        want = (it + 1) * (world + 1) / 2.0           # mean over ranks of (rank+1)*(it+1)
        got = p_arena[:n]
        bad += int((got - want).abs().max().item() > 1e-4 * want) + int(g_arena[:n].abs().max().item() != 0.0)
    torch.cuda.synchronize()
    t = torch.tensor([bad], device=dev, dtype=torch.float32)
    dist.all_reduce(t)
    ok = t.item() == 0
    results.append({"check": f"stress/skewed_arrival/{iters}_calls", "bad_calls": int(t.item()), "ok": ok})
    return ok


def stress_ll(comm, dev, world, rank, results, iters=96):
    """The flag-in-payload schedules have no barrier: consecutive calls are only kept apart by the
    two receive parities.  ``iters`` LL calls are queued back to back with NO host synchronisation,
    each rank stalling its stream by different random amounts in between, every call reducing into
    its own region; all regions are checked at the end."""
    import random
    if not comm.ll_capacity(C.ALGO_LL1):
        return True
    regions, n = 16, 8 * world * 515
    total = regions * n
    gbuf, pbuf = comm.alloc_pair(total * 4, total * 4)
    ctx = comm.make_ctx(gbuf, pbuf)
    g_arena, p_arena = gbuf.view(torch.float32, total), pbuf.view(torch.float32, total)
    hyper = make_hyper(dev, lr=1e-2, grad_scale=1.0)
    rng = random.Random(99 + 31 * rank)
    aux_out = torch.zeros(C.AUX_FLOATS, device=dev)
    aux_seen = []
    torch.cuda.synchronize()
    dist.barrier()
    for it in range(iters):
        r = it % regions
        g_arena[r * n:(r + 1) * n].fill_(float((rank + 1) * (it + 1)))
        comm.aux[:3].fill_(float(rank + it))
        if rng.random() < 0.5:
            torch.cuda._sleep(int(rng.random() * 2e6))
        comm.fused(ctx, algo=C.ALGO_LL1 if it % 3 else C.ALGO_LL2, opt=0, grad_dtype=torch.float32,
                   param_dtype=torch.float32, off=r * n, length=n, master=None, s1=None, s2=None,
                   state_off=0, hyper=hyper, zero_grads=True, aux_len=3, aux_out=aux_out)
        aux_seen.append(aux_out[:3].clone())
    torch.cuda.synchronize()
    bad = 0
    for it in range(iters - regions, iters):
        r = it % regions
        want = (it + 1) * world * (world + 1) / 2.0
        bad += int((p_arena[r * n:(r + 1) * n] - want).abs().max().item() > 1e-4 * want)
    for it, a in enumerate(aux_seen):
        want = sum(p + it for p in range(world))
        bad += int((a - want).abs().max().item() > 1e-3)
    bad += int(g_arena.abs().max().item() != 0.0) + int(comm.aux[:3].abs().max().item() != 0.0)
    t = torch.tensor([bad], device=dev, dtype=torch.float32)
    dist.all_reduce(t)
    ok = t.item() == 0
    results.append({"check": f"stress/ll_back_to_back/{iters}_calls", "bad": int(t.item()), "ok": ok})
    return ok


def time_fn(fn, iters, dev, graph=False):
    """Device time per call (CUDA events, max over ranks).  With ``graph`` the ``iters`` calls are
    captured into one CUDA graph and the replay is timed: for latency-bound sizes this removes the
    host's launch rate (a Python + ctypes call costs more than the kernel) from the number - what a
    training step sees, since the step is a CUDA graph too."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 1
    if graph:
        s = torch.cuda.Stream(dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            for i in range(3):
                fn(i)
        torch.cuda.current_stream(dev).wait_stream(s)
        torch.cuda.synchronize()
        dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(iters):
                fn(i)
        g.replay()
        reps = 5
        run = g.replay
    else:
        def run():
            for i in range(iters):
                fn(i)
    torch.cuda.synchronize()
    dist.barrier()
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / (iters * reps)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def sweep(comm, dev, world, rank, max_bytes, results):
    hyper = make_hyper(dev, lr=1e-3)
    cap = max(max_bytes, 1 << 20)
    gbuf, pbuf = comm.alloc_pair(cap, cap)
    ctx = comm.make_ctx(gbuf, pbuf)
    g_arena = gbuf.view(torch.float32, cap // 4)
    nccl_buf = torch.zeros(cap // 4, device=dev)
    has_mc = bool(gbuf.mc_ptr and pbuf.mc_ptr)
    size = 1 << 10
    while size <= max_bytes:
        n = max(size // 4, 8 * world)
        n = (n + 8 * world - 1) // (8 * world) * (8 * world)
        nbytes = n * 4
        slots = max(1, min(cap // nbytes, 8))          # rotate through the arena: defeats L2 reuse
        iters = 200 if nbytes <= (1 << 20) else (50 if nbytes <= (64 << 20) else 10)
        warm = 5
        row = {"bytes": nbytes, "world": world}

        def nccl(i):
            off = (i % slots) * n
            dist.all_reduce(nccl_buf[off:off + n])
        use_graph = nbytes <= (4 << 20)
        iters = min(iters, 50) if use_graph else iters
        for i in range(warm):
            nccl(i)
        ms = time_fn(nccl, iters, dev, graph=use_graph)
        row["nccl_us"] = ms * 1e3
        row["nccl_busbw_gbs"] = nbytes / (ms * 1e-3) * 2 * (world - 1) / world / 1e9
        algos = [("oneshot", C.ALGO_ONESHOT)] if nbytes <= (8 << 20) else []
        algos.append(("twoshot", C.ALGO_TWOSHOT))
        if has_mc:
            algos.append(("nvls", C.ALGO_NVLS))
        for name, algo in (("ll1", C.ALGO_LL1), ("ll2", C.ALGO_LL2)):
            if nbytes <= comm.ll_capacity(algo):
                algos.append((name, algo))
        for name, algo in algos:
            def ours(i, algo=algo):
                off = (i % slots) * n
                comm.fused(ctx, algo=algo, opt=0, grad_dtype=torch.float32,
                           param_dtype=torch.float32, off=off, length=n, master=None, s1=None,
                           s2=None, state_off=0, hyper=hyper, zero_grads=False)
            for i in range(warm):
                ours(i)
            ms = time_fn(ours, iters, dev, graph=use_graph)
            row[name + "_us"] = ms * 1e3
            row[name + "_busbw_gbs"] = nbytes / (ms * 1e-3) * 2 * (world - 1) / world / 1e9
        best = max((row.get(k + "_busbw_gbs", 0.0), k) for k in ("oneshot", "twoshot", "nvls", "ll1", "ll2"))
        row["best"] = best[1]
        row["best_vs_nccl"] = best[0] / row["nccl_busbw_gbs"] if row["nccl_busbw_gbs"] else None
        row["best_frac_of_900"] = best[0] / 900.0
        results.append(row)
        log(json.dumps(row))
        size *= 4 if size < (1 << 20) else 2 if size < (256 << 20) else 4


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check-only", action="store_true")
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--out", default="gpurun_out/allreduce_sweep.json")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    world, rank = dist.get_world_size(), dist.get_rank()
    comm = Communicator(dev, backend="p2p")
    out = {"world": world, "backend": comm.backend, "checks": [], "sweep": []}
    if comm.backend != "p2p":
        log("peer-memory backend unavailable:", comm.backend)
        out["error"] = "p2p backend unavailable"
    else:
        ok = check(comm, dev, world, rank, out["checks"])
        ok = stress(comm, dev, world, rank, out["checks"]) and ok
        ok = stress_ll(comm, dev, world, rank, out["checks"]) and ok
        out["checks_ok"] = bool(ok)
        for c in out["checks"]:
            log(json.dumps(c))
        log("CHECKS_OK" if ok else "CHECKS_FAILED")
        if not args.check_only:
            sweep(comm, dev, world, rank, args.max_mb << 20, out["sweep"])
    if rank == 0:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        with open(args.out.replace(".json", f"_n{world}.json"), "w") as f:
            json.dump(out, f, indent=1)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if out.get("checks_ok") else 1


if __name__ == "__main__":
    sys.exit(main())
