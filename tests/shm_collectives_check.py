"""Worker script (run under torch.distributed.run with the gloo backend): the shared-memory
collectives of cloud_b200/parallel/shm.py against torch.distributed as the oracle - fp32 and bf16,
one-shot and two-shot, ragged sizes, slices of a larger arena tensor, broadcast, the scalar side
channel - and, with SHM_CHECK_ABSENT_RANK, the watchdog."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200.parallel import shm  # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    arena = shm.ShmArena(8 << 20)
    absent = os.environ.get("SHM_CHECK_ABSENT_RANK")
    if absent is not None:
        arena.set_timeout(1.0)
        t = arena.tensor(1024, torch.float32)
        if rank == int(absent):
            time.sleep(3.0)                         # never joins the collective
        else:
            try:
                arena.all_reduce(t)
                raise SystemExit("the all-reduce returned although a rank never arrived")
            except shm.ShmError as e:
                print("WATCHDOG OK rank {}: {}".format(rank, e), flush=True)
        dist.barrier()
        return
    g = torch.Generator().manual_seed(100 + rank)
    big = arena.tensor(1 << 18, torch.float32)
    half = arena.tensor(70001, torch.bfloat16)
    for step, (numel, algo) in enumerate([(1, 0), (7, 1), (1000, 0), (1000, 1), (65537, None), (1 << 18, None)]):
        view = big[3:3 + numel] if numel + 3 <= big.numel() else big[:numel]
        view.copy_(torch.randn(view.numel(), generator=g))
        want = view.clone()
        dist.all_reduce(want)                      # oracle (sums in a different order: tolerance)
        arena.aux[0] += rank + 1.0
        arena.aux[1] += 0.5
        aux = arena.all_reduce(view, algo=algo, aux_len=2)
        assert torch.allclose(view, want, rtol=1e-5, atol=1e-5), (step, float((view - want).abs().max()))
        assert aux.tolist() == [world * (world + 1) / 2.0, 0.5 * world], aux
        # every rank holds bit-identical results (fixed summation order)
        mine = view.clone()
        root = mine.clone()
        dist.broadcast(root, src=0)
        assert torch.equal(mine, root), step
    for algo in (0, 1):
        half.copy_(torch.randn(half.numel(), generator=g).to(torch.bfloat16))
        want = half.float()
        dist.all_reduce(want)
        arena.all_reduce(half, scale=0.5, algo=algo)
        assert torch.allclose(half.float(), 0.5 * want, rtol=2e-2, atol=2e-2)
    big.fill_(float(rank))
    arena.broadcast(big[10:5000], root=world - 1)
    assert big[9] == rank and big[5000] == rank and bool((big[10:5000] == world - 1).all())
    assert torch.equal(arena.reduce_aux(2), torch.zeros(2))       # the side channel was cleared
    outside = torch.zeros(4)
    try:
        arena.all_reduce(outside)
        raise SystemExit("a tensor outside the arena was accepted")
    except shm.ShmError:
        pass
    arena.barrier()
    arena.close()
    if rank == 0:
        print("SHM OK world={}".format(world), flush=True)


if __name__ == "__main__":
    main()
