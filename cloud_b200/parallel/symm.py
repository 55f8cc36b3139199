"""Symmetric memory over NVLink: every rank maps every peer's arena (and an NVLS multicast view).

Native part: ``csrc/runtime/symm_arena.cc`` (CUDA VMM + multicast driver calls).  This module does
the control plane only: it swaps the POSIX file descriptors that name the allocations between the
ranks of one box with ``SCM_RIGHTS`` messages over AF_UNIX sockets, using ``torch.distributed``
just to agree on socket paths and to barrier.

No counterpart exists in the reference (it has no communication layer);
this is the memory substrate for the K9/K10 kernels.
"""
from __future__ import annotations

import ctypes
import os
import socket
import struct
import tempfile
import uuid

import torch
import torch.distributed as dist

from ..ops import _lib


class SymmUnavailable(RuntimeError):
    pass


def _rt():
    lib = _lib.runtime()
    if not getattr(lib, "_symm_declared", False):
        u64p = ctypes.POINTER(ctypes.c_uint64)
        ip = ctypes.POINTER(ctypes.c_int)
        szp = ctypes.POINTER(ctypes.c_size_t)
        lib.cb2rt_last_error.restype = ctypes.c_char_p
        lib.cb2rt_symm_capabilities.argtypes = [ctypes.c_int, ip, ip]
        lib.cb2rt_symm_granularity.argtypes = [ctypes.c_int, ctypes.c_int, szp]
        lib.cb2rt_symm_create.argtypes = [ctypes.c_int, ctypes.c_size_t, u64p, ip]
        lib.cb2rt_symm_import.argtypes = [ctypes.c_int, u64p]
        lib.cb2rt_symm_map.argtypes = [ctypes.c_uint64, ctypes.c_int, ctypes.c_size_t,
                                       ctypes.c_size_t, u64p]
        lib.cb2rt_symm_unmap.argtypes = [ctypes.c_uint64, ctypes.c_size_t]
        lib.cb2rt_symm_release.argtypes = [ctypes.c_uint64]
        lib.cb2rt_mc_create.argtypes = [ctypes.c_int, ctypes.c_size_t, u64p, ip]
        lib.cb2rt_mc_add_device.argtypes = [ctypes.c_uint64, ctypes.c_int]
        lib.cb2rt_mc_bind.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_size_t]
        lib.cb2rt_close_fd.argtypes = [ctypes.c_int]
        lib._symm_declared = True
    return lib


def _ok(rc: int, what: str) -> None:
    if rc != 0:
        raise SymmUnavailable(f"{what}: {_rt().cb2rt_last_error().decode(errors='replace')}")


def capabilities(device_index: int) -> dict:
    """{'vmm': bool, 'multicast': bool} for a CUDA device (False/False without a driver)."""
    vmm, mc = ctypes.c_int(0), ctypes.c_int(0)
    try:
        rc = _rt().cb2rt_symm_capabilities(device_index, ctypes.byref(vmm), ctypes.byref(mc))
    except Exception:
        return {"vmm": False, "multicast": False}
    if rc != 0:
        return {"vmm": False, "multicast": False}
    return {"vmm": bool(vmm.value), "multicast": bool(mc.value)}


class _RawCudaBuffer:
    """Exposes a raw device pointer through ``__cuda_array_interface__`` so torch can wrap it."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None}


def _wrap(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCudaBuffer(ptr, nbytes), device=device)


class FdExchange:
    """All-to-all exchange of file descriptors between the ranks of one host."""

    def __init__(self, group=None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        token = [uuid.uuid4().hex[:12] if self.rank == 0 else None]
        dist.broadcast_object_list(token, src=dist.get_global_rank(group, 0) if group else 0,
                                   group=group)
        self.base = os.path.join(tempfile.gettempdir(), f"cb2symm_{token[0]}")
        self.path = f"{self.base}_{self.rank}.sock"
        self.server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        if os.path.exists(self.path):
            os.unlink(self.path)
        self.server.bind(self.path)
        self.server.listen(self.world + 2)
        dist.barrier(group=group)

    def all_to_all(self, fd: int) -> list[int]:
        """Every rank contributes one fd; returns the world's fds (own slot = the given fd)."""
        for p in range(self.world):
            if p == self.rank:
                continue
            s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            s.connect(f"{self.base}_{p}.sock")
            socket.send_fds(s, [struct.pack("i", self.rank)], [fd])
            s.close()
        out = [-1] * self.world
        out[self.rank] = fd
        for _ in range(self.world - 1):
            conn, _addr = self.server.accept()
            msg, fds, _flags, _a = socket.recv_fds(conn, 4, 1)
            conn.close()
            out[struct.unpack("i", msg)[0]] = fds[0]
        dist.barrier(group=self.group)
        return out

    def one_to_all(self, fd: int | None, root: int = 0) -> int:
        """`root` sends one fd to everybody; returns it (root returns its own)."""
        if self.rank == root:
            for p in range(self.world):
                if p == root:
                    continue
                s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                s.connect(f"{self.base}_{p}.sock")
                socket.send_fds(s, [struct.pack("i", self.rank)], [fd])
                s.close()
            got = fd
        else:
            conn, _addr = self.server.accept()
            _msg, fds, _flags, _a = socket.recv_fds(conn, 4, 1)
            conn.close()
            got = fds[0]
        dist.barrier(group=self.group)
        return got

    def close(self):
        try:
            self.server.close()
            os.unlink(self.path)
        except OSError:
            pass


class SymmetricBuffer:
    """One symmetric allocation: ``local`` (uint8 tensor), ``peer_ptrs`` (my mapping of every
    rank's copy), ``mc_ptr`` (multicast mapping or 0)."""

    def __init__(self, nbytes: int, device: torch.device, xchg: FdExchange | None,
                 want_multicast: bool = True):
        self.device = device
        self.dev_index = device.index if device.index is not None else torch.cuda.current_device()
        self.rank = xchg.rank if xchg else 0
        self.world = xchg.world if xchg else 1
        rt = _rt()
        gran = ctypes.c_size_t(0)
        _ok(rt.cb2rt_symm_granularity(self.dev_index, self.world, ctypes.byref(gran)), "granularity")
        g = max(gran.value, 2 << 20)
        self.nbytes = ((nbytes + g - 1) // g) * g
        self.align = g
        handle, fd = ctypes.c_uint64(0), ctypes.c_int(-1)
        _ok(rt.cb2rt_symm_create(self.dev_index, self.nbytes, ctypes.byref(handle), ctypes.byref(fd)),
            "create")
        self.handle = handle.value
        self._handles = [self.handle]
        self.peer_ptrs = [0] * self.world
        ptr = ctypes.c_uint64(0)
        _ok(rt.cb2rt_symm_map(self.handle, self.dev_index, self.nbytes, g, ctypes.byref(ptr)), "map")
        self.peer_ptrs[self.rank] = ptr.value
        self.local = _wrap(ptr.value, self.nbytes, device)
        self.local.zero_()
        torch.cuda.synchronize(device)
        self.mc_ptr = 0
        if self.world > 1:
            fds = xchg.all_to_all(fd.value)
            for p, pfd in enumerate(fds):
                if p == self.rank:
                    continue
                ph = ctypes.c_uint64(0)
                _ok(rt.cb2rt_symm_import(pfd, ctypes.byref(ph)), f"import from rank {p}")
                rt.cb2rt_close_fd(pfd)
                self._handles.append(ph.value)
                pp = ctypes.c_uint64(0)
                _ok(rt.cb2rt_symm_map(ph.value, self.dev_index, self.nbytes, g, ctypes.byref(pp)),
                    f"map rank {p}")
                self.peer_ptrs[p] = pp.value
            if want_multicast:
                try:
                    self._setup_multicast(xchg)
                except SymmUnavailable:
                    self.mc_ptr = 0
        rt.cb2rt_close_fd(fd.value)

    def _setup_multicast(self, xchg: FdExchange) -> None:
        rt = _rt()
        caps = capabilities(self.dev_index)
        flags = [None] * self.world
        dist.all_gather_object(flags, bool(caps["multicast"]), group=xchg.group)
        if not all(flags):
            return
        mc, mfd = ctypes.c_uint64(0), ctypes.c_int(-1)
        ok = True
        if self.rank == 0:
            ok = rt.cb2rt_mc_create(self.world, self.nbytes, ctypes.byref(mc), ctypes.byref(mfd)) == 0
        oks = [None] * self.world
        dist.all_gather_object(oks, ok, group=xchg.group)
        if not all(oks):
            return
        got = xchg.one_to_all(mfd.value if self.rank == 0 else None, root=0)
        if self.rank != 0:
            _ok(rt.cb2rt_symm_import(got, ctypes.byref(mc)), "import multicast")
        rt.cb2rt_close_fd(got)
        ok = rt.cb2rt_mc_add_device(mc.value, self.dev_index) == 0
        dist.all_gather_object(oks, ok, group=xchg.group)   # also the "all devices added" barrier
        if not all(oks):
            return
        ok = rt.cb2rt_mc_bind(mc.value, self.handle, self.nbytes) == 0
        dist.all_gather_object(oks, ok, group=xchg.group)
        if not all(oks):
            return
        mp = ctypes.c_uint64(0)
        ok = rt.cb2rt_symm_map(mc.value, self.dev_index, self.nbytes, self.align, ctypes.byref(mp)) == 0
        dist.all_gather_object(oks, ok, group=xchg.group)
        if all(oks):
            self.mc_ptr = mp.value
            self._handles.append(mc.value)

    def view(self, dtype: torch.dtype, numel: int | None = None, byte_offset: int = 0) -> torch.Tensor:
        t = self.local[byte_offset:]
        item = torch.empty((), dtype=dtype).element_size()
        n = numel if numel is not None else t.numel() // item
        return t[: n * item].view(dtype)
