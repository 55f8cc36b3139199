"""Communicator: the data-parallel engine's view of "the other replicas".

Four back ends behind one interface:

* ``p2p``   - hand-written sm_100a kernels over symmetric NVLink memory
              (``csrc/kernels/comm.cu``: one-shot / two-shot / NVLS, fused with the optimizer).
              This is the product path on a multi-GPU B200 box.
* ``local`` - world size 1: the same fused kernels with no peers (CUDA) or plain PyTorch (CPU).
* ``shm``   - CPU ranks of one box: the one-shot / two-shot schedules and the epoch-flag protocol of
              the kernels over a POSIX shared-memory segment (``csrc/runtime/shm_comm.cc``);
              gradients live in the segment, so the all-reduce is in place.  Default for CPU ranks.
* ``dist``  - ``torch.distributed`` all-reduce (NCCL on GPUs, gloo on CPUs) followed by the flat
              optimizer kernel.  Baseline / oracle / CPU test path, and the fallback when peer
              mapping is impossible; never chosen silently on a box where p2p works.

Reference parity: what ``tf.distribute`` cross-device ops did for MirroredStrategy /
MultiWorkerMirroredStrategy (core/preprocess.py:139-143) - see SURVEY.md sections 2.4 and 5.8.
"""
from __future__ import annotations

import ctypes
import os
import warnings

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import CommCtx

MAX_RANKS = 8
COMM_MAX_BLOCKS = 128
AUX_FLOATS = 256
PAD_BYTES = COMM_MAX_BLOCKS * MAX_RANKS * 4 + AUX_FLOATS * 4

ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS, ALGO_LL1, ALGO_LL2 = 0, 1, 2, 3, 4
ALGO_NAMES = {"oneshot": ALGO_ONESHOT, "twoshot": ALGO_TWOSHOT, "nvls": ALGO_NVLS, "ll1": ALGO_LL1,
              "ll2": ALGO_LL2}
ALGO_LABELS = {v: k for k, v in ALGO_NAMES.items()}
SHARDED_ALGOS = (ALGO_TWOSHOT, ALGO_NVLS, ALGO_LL2)       # optimizer state cut across the ranks
# gradient bytes one LL slot can take: the largest LL1 bucket, and 1/world of the largest LL2 bucket
LL_SLOT_PAYLOAD = int(os.environ.get("TFC_LL_SLOT_KB", "512")) << 10

_launches = 0


def launches() -> int:
    return _launches


class Communicator:
    def __init__(self, device: torch.device, group=None, backend: str | None = None,
                 single: bool = False):
        self.device = device
        self.group = group
        # `single`: a one-replica strategy inside a process that also belongs to a process group
        self.initialized = (not single) and dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if self.initialized else 0
        self.world = dist.get_world_size(group) if self.initialized else 1
        env = os.environ.get("TFC_COMM_BACKEND")
        backend = backend or env
        if backend is None:
            if self.world == 1:
                backend = "local"
            elif device.type == "cuda":
                backend = "p2p"
            else:
                backend = "shm"
        self.backend = backend
        self.shm = None            # ShmArena (CPU ranks)
        self.xchg = None
        self.pad = None            # SymmetricBuffer (flags + aux)
        self.epochs = None
        self.aux = None            # fp32 view of the aux scalars (local)
        self.multicast = False
        self.ll = None
        self.ll_state = None
        if device.type == "cuda" and backend in ("p2p", "local"):
            self._init_symm()
        if backend == "shm":
            self._init_shm()
        if self.aux is None:
            self.aux = torch.zeros(AUX_FLOATS, dtype=torch.float32, device=device)

    # -- set-up -------------------------------------------------------------------------------
    def _init_symm(self) -> None:
        from . import symm
        if self.world > MAX_RANKS:
            raise RuntimeError(f"p2p backend supports up to {MAX_RANKS} ranks per box")
        try:
            if self.world > 1:
                self.xchg = symm.FdExchange(self.group)
            self.pad = symm.SymmetricBuffer(PAD_BYTES, self.device, self.xchg, want_multicast=False)
        except (symm.SymmUnavailable, OSError) as e:
            if self.world > 1:
                oks = [None] * self.world
                dist.all_gather_object(oks, False, group=self.group)
                warnings.warn(f"cloud_b200: peer-memory arena unavailable ({e}); using "
                              f"torch.distributed all-reduce instead")
                self.backend = "dist"
            else:
                self.backend = "local"
            self.pad = None
            return
        if self.world > 1:
            oks = [None] * self.world
            dist.all_gather_object(oks, True, group=self.group)
            if not all(oks):
                self.backend = "dist"
                self.pad = None
                return
        self.epochs = torch.zeros(COMM_MAX_BLOCKS, dtype=torch.int32, device=self.device)
        self.aux = self.pad.view(torch.float32, AUX_FLOATS, COMM_MAX_BLOCKS * MAX_RANKS * 4)
        # receive buffers of the flag-in-payload (LL) schedules, shared by every context of this
        # communicator, and their call counter
        self.ll = None
        self.ll_state = torch.zeros(2, dtype=torch.int32, device=self.device)
        if self.world in (2, 4, 8) and os.environ.get("TFC_DISABLE_LL") != "1":
            lib = _lib.kernels()
            lib.cb2_comm_ll_bytes.restype = ctypes.c_size_t
            lib.cb2_comm_ll_bytes.argtypes = [ctypes.c_int, ctypes.c_size_t]
            nbytes = int(lib.cb2_comm_ll_bytes(self.world, LL_SLOT_PAYLOAD))
            try:
                self.ll = symm.SymmetricBuffer(nbytes, self.device, self.xchg, want_multicast=False)
                ok = True
            except (symm.SymmUnavailable, OSError):
                ok = False
            oks = [None] * self.world
            dist.all_gather_object(oks, ok, group=self.group)
            if not all(oks):
                self.ll = None

    def _init_shm(self) -> None:
        from . import shm
        if self.device.type != "cpu" or self.world == 1:
            self.backend = "dist" if self.world > 1 else "local"
            return
        try:
            self.shm = shm.ShmArena(1 << 20, self.group)        # signal pad + scalars + small tensors
        except (shm.ShmError, _lib.KernelsUnavailable, OSError) as e:
            warnings.warn(f"cloud_b200: shared-memory arena unavailable ({e}); using "
                          f"torch.distributed all-reduce instead")
            self.backend = "dist"
            self.shm = None
            return
        self._shm_arenas = [self.shm]
        self.aux = self.shm.aux

    @property
    def uses_peer_memory(self) -> bool:
        return self.pad is not None

    def alloc_flat(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """A zero-filled flat tensor for gradients.  CPU ranks with the shared-memory back end get it
        inside a segment of exactly that size (reserved up front), so ``all_reduce_sum`` on it - or
        on a slice of it - is in place.  Collective there: every rank calls it in the same order."""
        if self.shm is not None:
            from . import shm
            nbytes = numel * torch.empty((), dtype=dtype).element_size()
            cap = int(float(os.environ.get("TFC_SHM_ARENA_MB", "4096")) * (1 << 20))
            err = None
            if nbytes <= cap:
                try:
                    arena = shm.ShmArena(nbytes + 4096, self.group)
                    self._shm_arenas.append(arena)
                    return arena.tensor(numel, dtype)
                except shm.ShmError as e:          # raised on every rank or on none (see ShmArena)
                    err = e
            warnings.warn(f"cloud_b200: {nbytes >> 20} MiB of gradients do not go into shared memory "
                          f"({err or 'above TFC_SHM_ARENA_MB'}); this group uses torch.distributed")
        return torch.zeros(numel, dtype=dtype, device=self.device)

    def _shm_arena_of(self, t: torch.Tensor):
        if self.shm is None or t.dtype not in (torch.float32, torch.bfloat16):
            return None
        for arena in self._shm_arenas:
            if arena.offset_of(t) is not None:
                return arena
        return None

    def alloc_pair(self, nbytes_grad: int, nbytes_param: int):
        """Symmetric (grad, param) arenas for one parameter group; plain tensors otherwise."""
        from . import symm
        if self.pad is not None:
            want_mc = self.world > 1 and os.environ.get("TFC_DISABLE_NVLS") != "1"
            g = symm.SymmetricBuffer(nbytes_grad, self.device, self.xchg, want_multicast=want_mc)
            p = symm.SymmetricBuffer(nbytes_param, self.device, self.xchg, want_multicast=want_mc)
            return g, p
        return None, None

    def make_ctx(self, grad_buf, param_buf) -> CommCtx:
        ctx = CommCtx()
        ctx.rank, ctx.world = self.rank, self.world
        for i in range(MAX_RANKS):
            ctx.grad_ptrs[i] = grad_buf.peer_ptrs[i] if i < self.world else None
            ctx.param_ptrs[i] = param_buf.peer_ptrs[i] if i < self.world else None
            ctx.pad_ptrs[i] = self.pad.peer_ptrs[i] if i < self.world else None
        ctx.mc_grad = grad_buf.mc_ptr or None
        ctx.mc_param = param_buf.mc_ptr or None
        ctx.epochs = self.epochs.data_ptr()
        for i in range(MAX_RANKS):
            ctx.ll_ptrs[i] = self.ll.peer_ptrs[i] if (self.ll is not None and i < self.world) else None
        ctx.ll_bytes = self.ll.nbytes if self.ll is not None else 0
        ctx.ll_state = self.ll_state.data_ptr() if self.ll_state is not None else None
        return ctx

    def ll_capacity(self, algo: int) -> int:
        """Largest bucket (bytes) the LL buffers take under ``algo`` (0: LL unavailable)."""
        if self.ll is None:
            return 0
        # mirrors ll_slot_chunks() in csrc/kernels/comm.cu
        aux = self.world * AUX_FLOATS * 8
        chunks = (self.ll.nbytes // 2 - aux) // 16 // (4 * self.world)
        return chunks * 16 if algo == ALGO_LL1 else chunks * 16 * self.world

    # -- kernels ------------------------------------------------------------------------------
    def fused(self, ctx: CommCtx, *, algo: int, opt: int, grad_dtype: torch.dtype,
              param_dtype: torch.dtype, off: int, length: int, master, s1, s2, state_off: int,
              hyper: torch.Tensor, zero_grads: bool, aux_len: int = 0, aux_out=None,
              blocks: int = 0, stream=None) -> None:
        global _launches
        st = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        rc = _lib.kernels().cb2_comm_fused(
            ctypes.byref(ctx), algo, opt, 1 if grad_dtype == torch.bfloat16 else 0,
            1 if param_dtype == torch.bfloat16 else 0, off, length,
            None if master is None else master.data_ptr(),
            None if s1 is None else s1.data_ptr(), None if s2 is None else s2.data_ptr(),
            state_off, hyper.data_ptr(), int(zero_grads), aux_len,
            None if aux_out is None else aux_out.data_ptr(), blocks, st)
        _lib.check(rc, "comm_fused")
        _launches += 1

    def broadcast_arena(self, ctx: CommCtx, dtype: torch.dtype, off: int, length: int,
                        root: int = 0) -> None:
        global _launches
        rc = _lib.kernels().cb2_comm_broadcast(
            ctypes.byref(ctx), root, 1 if dtype == torch.bfloat16 else 0, off, length,
            1 if ctx.mc_param else 0, 0, torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(rc, "comm_broadcast")
        _launches += 1

    def device_barrier(self, ctx: CommCtx) -> None:
        global _launches
        rc = _lib.kernels().cb2_comm_barrier(ctypes.byref(ctx),
                                             torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(rc, "comm_barrier")
        _launches += 1

    # -- torch.distributed fallbacks ----------------------------------------------------------
    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        arena = self._shm_arena_of(t) if self.world > 1 else None
        if arena is not None:
            arena.all_reduce(t)
            return t
        if self.world > 1:
            if t.dtype == torch.bfloat16 and t.device.type == "cpu":
                f = t.float()
                dist.all_reduce(f, group=self.group)
                t.copy_(f.to(t.dtype))
            else:
                dist.all_reduce(t, group=self.group)
        return t

    def broadcast(self, t: torch.Tensor, root: int = 0) -> torch.Tensor:
        if self.world > 1:
            src = dist.get_global_rank(self.group, root) if self.group else root
            dist.broadcast(t, src=src, group=self.group)
        return t

    def barrier(self) -> None:
        if self.world > 1:
            dist.barrier(group=self.group)
