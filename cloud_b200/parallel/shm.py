"""Shared-memory collectives for the CPU ranks of one box (``csrc/runtime/shm_comm.cc``).

The host-memory twin of the peer-memory kernels: every rank owns one arena of a POSIX shared-memory
segment that all ranks map, gradients are produced straight into it (``ShmArena.tensor`` hands out
views), and ``all_reduce`` runs the one-shot or two-shot schedule with the same epoch-flag protocol
as ``csrc/kernels/comm.cu`` - so ``MirroredStrategy`` on CPU presets exercises the product's
schedules rather than a library ring.  ``torch.distributed`` (gloo) stays the control plane: it
carries the segment name and the "everybody has mapped it" barrier, like the fd exchange of the
GPU arena (``symm.py``).  Reference parity: what TensorFlow's CPU collectives did underneath
``tf.distribute`` (core/preprocess.py:137-143).
"""
from __future__ import annotations

import ctypes
import os
import uuid

import torch
import torch.distributed as dist

from ..ops import _lib

ALGO_ONESHOT, ALGO_TWOSHOT = 0, 1
ONESHOT_MAX_BYTES = 16 << 10           # measured cross-over (bench/shm_allreduce.py, 2 and 4 ranks)


class ShmError(RuntimeError):
    pass


_declared = False


def _rt():
    global _declared
    lib = _lib.runtime()
    if not _declared:
        vp, i, sz, d, f = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_double, ctypes.c_float
        lib.cb2rt_shm_open.restype = vp
        lib.cb2rt_shm_open.argtypes = [ctypes.c_char_p, i, i, sz, i, d]
        lib.cb2rt_shm_unlink.argtypes = [ctypes.c_char_p]
        lib.cb2rt_shm_close.argtypes = [vp]
        lib.cb2rt_shm_arena.restype = vp
        lib.cb2rt_shm_arena.argtypes = [vp, i]
        lib.cb2rt_shm_aux.restype = vp
        lib.cb2rt_shm_aux.argtypes = [vp]
        lib.cb2rt_shm_aux_floats.restype = i
        lib.cb2rt_shm_barrier.argtypes = [vp]
        lib.cb2rt_shm_allreduce.argtypes = [vp, i, i, sz, sz, f, i, vp]
        lib.cb2rt_shm_broadcast.argtypes = [vp, i, sz, sz]
        lib.cb2rt_shm_set_timeout.argtypes = [vp, d]
        lib.cb2rt_last_error.restype = ctypes.c_char_p
        _declared = True
    return lib


def _check(rc, what):
    if rc != 0:
        raise ShmError("{} failed: {}".format(what, _rt().cb2rt_last_error().decode(errors="replace")))


class ShmArena:
    """One rank's handle on the segment: ``arena_bytes`` of symmetric memory per rank."""

    def __init__(self, arena_bytes: int, group=None, timeout: float | None = None):
        lib = _rt()
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.arena_bytes = int(arena_bytes)
        timeout = float(os.environ.get("TFC_SHM_TIMEOUT", "120")) if timeout is None else timeout
        names = [None]
        if self.rank == 0:
            names[0] = "/cb2_{}_{}".format(os.getpid(), uuid.uuid4().hex[:12])
            self._h = lib.cb2rt_shm_open(names[0].encode(), 0, self.world, self.arena_bytes, 1, timeout)
            if not self._h:
                names[0] = "!" + lib.cb2rt_last_error().decode(errors="replace")
        src = dist.get_global_rank(group, 0) if group is not None else 0
        dist.broadcast_object_list(names, src=src, group=group)
        if names[0].startswith("!"):
            raise ShmError("creating the shared-memory segment failed: " + names[0][1:])
        self.name = names[0]
        if self.rank != 0:
            self._h = lib.cb2rt_shm_open(self.name.encode(), self.rank, self.world, self.arena_bytes, 0, timeout)
        oks = [None] * self.world
        dist.all_gather_object(oks, bool(self._h), group=group)
        if self.rank == 0:
            lib.cb2rt_shm_unlink(self.name.encode())      # everybody holds a mapping: no name, no leak
        if not all(oks):
            if self._h:
                lib.cb2rt_shm_close(self._h)
                self._h = None
            raise ShmError("rank(s) {} could not map the shared-memory segment".format(
                [r for r, ok in enumerate(oks) if not ok]))
        self._base = lib.cb2rt_shm_arena(self._h, self.rank)
        self._cursor = 0
        n_aux = lib.cb2rt_shm_aux_floats()
        self.aux = self._view(lib.cb2rt_shm_aux(self._h), n_aux * 4, torch.float32, n_aux)
        self._aux_out = torch.zeros(n_aux, dtype=torch.float32)

    # -- memory ---------------------------------------------------------------------------------
    def _view(self, addr, nbytes, dtype, numel):
        buf = (ctypes.c_char * nbytes).from_address(addr)
        t = torch.frombuffer(buf, dtype=dtype, count=numel)
        t._cb2_keepalive = (buf, self)
        return t

    def tensor(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """A zero-filled flat tensor inside this rank's arena (same offset on every rank when every
        rank makes the same sequence of calls)."""
        item = torch.empty((), dtype=dtype).element_size()
        off = (self._cursor + 63) // 64 * 64
        nbytes = int(numel) * item
        if off + nbytes > self.arena_bytes:
            raise ShmError("arena of {} bytes is full ({} + {} requested)".format(self.arena_bytes, off, nbytes))
        self._cursor = off + nbytes
        t = self._view(self._base + off, max(nbytes, 1), dtype, int(numel)) if numel else torch.zeros(0, dtype=dtype)
        t.zero_()
        return t

    def offset_of(self, t: torch.Tensor):
        """Byte offset of ``t`` inside this rank's arena, or None if it lives elsewhere."""
        if t.device.type != "cpu" or not t.is_contiguous():
            return None
        off = t.data_ptr() - self._base
        return off if 0 <= off and off + t.numel() * t.element_size() <= self.arena_bytes else None

    @property
    def free_bytes(self) -> int:
        return self.arena_bytes - self._cursor

    # -- collectives ------------------------------------------------------------------------------
    def all_reduce(self, t: torch.Tensor, scale: float = 1.0, algo: int | None = None, aux_len: int = 0):
        """In-place sum over ranks of a tensor that lives in the arena (see ``tensor``).  With
        ``aux_len`` the first scalars of ``aux`` are summed too and returned."""
        off = self.offset_of(t)
        if off is None:
            raise ShmError("all_reduce needs a tensor allocated from this arena")
        if t.dtype not in (torch.float32, torch.bfloat16):
            raise ShmError("shared-memory all-reduce supports float32 and bfloat16, got {}".format(t.dtype))
        item = t.element_size()
        if algo is None:
            algo = ALGO_ONESHOT if t.numel() * item <= ONESHOT_MAX_BYTES else ALGO_TWOSHOT
        rc = _rt().cb2rt_shm_allreduce(self._h, algo, 1 if t.dtype == torch.bfloat16 else 0, off // item,
                                       t.numel(), float(scale), aux_len, self._aux_out.data_ptr())
        _check(rc, "shared-memory all-reduce")
        return self._aux_out[:aux_len] if aux_len else None

    def reduce_aux(self, n: int) -> torch.Tensor:
        """Sum over ranks of the first ``n`` scalars of ``aux`` (which is cleared)."""
        rc = _rt().cb2rt_shm_allreduce(self._h, ALGO_ONESHOT, 0, 0, 0, 1.0, int(n), self._aux_out.data_ptr())
        _check(rc, "shared-memory scalar reduction")
        return self._aux_out[:n]

    def broadcast(self, t: torch.Tensor, root: int = 0) -> None:
        off = self.offset_of(t)
        if off is None:
            raise ShmError("broadcast needs a tensor allocated from this arena")
        _check(_rt().cb2rt_shm_broadcast(self._h, root, off, t.numel() * t.element_size()),
               "shared-memory broadcast")

    def barrier(self) -> None:
        _check(_rt().cb2rt_shm_barrier(self._h), "shared-memory barrier")

    def set_timeout(self, seconds: float) -> None:
        _rt().cb2rt_shm_set_timeout(self._h, float(seconds))

    def close(self) -> None:
        if getattr(self, "_h", None):
            _rt().cb2rt_shm_close(self._h)
            self._h = None
