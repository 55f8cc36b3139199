"""DPEngine - the data-parallel training engine behind every strategy.

What TensorFlow did for the reference under ``MirroredStrategy`` (mirrored variables, gradient
all-reduce, per-variable optimizer ops: SURVEY.md section 3.2) is re-designed here for B200:

* parameters and gradients live in **flat arenas** from birth (bf16 arena for matrix/conv weights
  with fp32 master copies, fp32 arena for biases / BatchNorm scales); with peers, the arenas are
  symmetric NVLink memory, so the collective needs no pack/unpack,
* arenas are cut into **buckets in reverse creation order**; an autograd post-accumulate hook
  launches a bucket's kernel on a side stream as soon as its last gradient is written, overlapping
  communication + optimizer with the rest of backward,
* one kernel per bucket does **all-reduce + scale + Adam/SGD/RMSprop + parameter publication +
  gradient zeroing** (``cb2_comm_fused``); optimizer state is sharded across ranks (ZeRO-1 style)
  in the two-shot / NVLS schedules and replicated in the one-shot schedule,
* step counter, bias corrections and learning rate sit in device memory, so the whole step can be
  captured in a CUDA graph.

The ``dist`` path (torch.distributed all-reduce + flat optimizer kernel) is the baseline/oracle and
the CPU (gloo) implementation of identical semantics.
"""
from __future__ import annotations

import os
import time
import weakref
from dataclasses import dataclass, field

import torch

from .. import ops
from ..ops.optim import OPT_CODES, make_hyper, optim_flat, optim_tick
from . import comm as comm_mod
from .comm import Communicator


@dataclass
class OptimizerSpec:
    """Optimizer description handed to the engine (built by cloud_b200.keras.optimizers)."""
    name: str = "adam"            # adam | adam_torch | adamw | sgd | rmsprop
    lr: float = 1e-3
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-7
    weight_decay: float = 0.0
    momentum: float = 0.0         # sgd momentum, or rmsprop rho
    nesterov: bool = False
    rms_momentum: float = 0.0

    @property
    def code(self) -> int:
        return OPT_CODES[self.name]


@dataclass
class _Bucket:
    off: int
    length: int
    n_params: int
    ready: int = 0
    launched: bool = False
    algo: int = 0            # collective schedule of this bucket (comm.ALGO_*)
    sharded: bool = False    # optimizer state of this bucket cut across the ranks
    s_off: int = 0           # where this bucket's part starts in this rank's state arrays
    s_len: int = 0


@dataclass
class _Group:
    dtype: torch.dtype
    params: list = field(default_factory=list)     # in arena order (reverse creation order)
    offsets: list = field(default_factory=list)
    total: int = 0
    param_arena: torch.Tensor | None = None
    grad_arena: torch.Tensor | None = None
    buckets: list = field(default_factory=list)
    master: torch.Tensor | None = None
    s1: torch.Tensor | None = None
    s2: torch.Tensor | None = None
    ctx: object = None
    bufs: tuple = ()
    algo: int = 0
    sharded: bool = False
    master_buf: object = None


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


_LIVE: "weakref.WeakSet" = weakref.WeakSet()


def live_engines() -> list:
    """Engines of this process (custom training loops arm them when a GradientTape opens)."""
    return list(_LIVE)


class DPEngine:
    def __init__(self, named_params, optimizer: OptimizerSpec, communicator: Communicator,
                 *, compute_dtype: torch.dtype | None = None, bucket_mb: float | None = None,
                 algo: str | None = None, grad_scale: float | None = None):
        self.comm = communicator
        self.device = communicator.device
        self.world = communicator.world
        self.rank = communicator.rank
        self.opt = optimizer
        self.is_cuda = self.device.type == "cuda"
        if compute_dtype is None:
            compute_dtype = torch.bfloat16 if self.is_cuda else torch.float32
        self.compute_dtype = compute_dtype
        self.bucket_elems = int(float(bucket_mb if bucket_mb is not None else
                                      os.environ.get("TFC_BUCKET_MB", "32")) * (1 << 20))
        self.algo_name = algo or os.environ.get("TFC_ALLREDUCE_ALGO", "auto")
        # gradients are sums over replicas of per-replica losses already divided by the GLOBAL
        # batch (Keras/tf.distribute convention) => no extra 1/world scale by default
        self.hyper = make_hyper(self.device, lr=optimizer.lr, beta1=optimizer.beta1,
                                beta2=optimizer.beta2, eps=optimizer.eps,
                                weight_decay=optimizer.weight_decay,
                                grad_scale=1.0 if grad_scale is None else grad_scale,
                                momentum=optimizer.momentum, nesterov=optimizer.nesterov,
                                rms_momentum=optimizer.rms_momentum)
        self._lr_host = None
        self.comm_stream = torch.cuda.Stream(self.device) if self.is_cuda else None
        self.groups: list[_Group] = []
        self.steps = 0
        self.aux_out = torch.zeros(comm_mod.AUX_FLOATS, dtype=torch.float32, device=self.device)
        self._hook_handles = []
        self.allreduce_host_us: list = []      # drained by the front end into the metrics registry
        self._in_step = False
        self._aux_pending = 0
        self._aux_done = 0
        self._direct = communicator.device.type == "cuda"
        self._written = set()
        self._bucket_events = []
        self._time_buckets = communicator.device.type == "cuda" and os.environ.get("TFC_TIME_BUCKETS", "1") != "0"
        self._master_exact = set()
        self._fault = None
        fault = os.environ.get("TFC_FAULT")
        if fault and os.environ.get("TFC_RESTART_COUNT", "0") == "0":   # injected once, not again after a restart
            try:
                r, s = fault.split(":")
                self._fault = (int(r), int(s))
            except ValueError:
                raise ValueError("TFC_FAULT must look like '<rank>:<step>', got {!r}".format(fault))
        self._build(list(named_params))
        _LIVE.add(self)

    @property
    def in_step(self) -> bool:
        return self._in_step

    # ------------------------------------------------------------------------------------------
    def _build(self, named_params) -> None:
        params = [(n, p) for n, p in named_params if p.requires_grad]
        lowp, f32 = _Group(self.compute_dtype), _Group(torch.float32)
        for name, p in reversed(params):
            g = lowp if (p.dim() >= 2 and self.compute_dtype != torch.float32) else f32
            g.params.append((name, p))
        self.groups = [g for g in (lowp, f32) if g.params]
        align = 8 * self.world
        for g in self.groups:
            itemsize = torch.empty((), dtype=g.dtype).element_size()
            sizes = [_round_up(p.numel(), 8) for _n, p in g.params]
            off = 0
            for lo, hi in self._partition(sizes, itemsize):
                b_start = off
                for i in range(lo, hi):
                    g.offsets.append(off)
                    off += sizes[i]
                off = _round_up(off - b_start, align) + b_start
                g.buckets.append(_Bucket(b_start, off - b_start, hi - lo))
            g.total = off
            self._alloc_group(g, itemsize)
            self._adopt_params(g)
        self.sync_from_root()
        self._install_hooks()

    def _partition(self, sizes, itemsize):
        """Cut the arena (parameters in the order their gradients become ready) into buckets.
        The bucket that is ready LAST - the first layers of the model - cannot hide behind any
        backward work, so it is kept small (``TFC_TAIL_BUCKET_MB``, 4); towards the front of the
        arena, where all of backward is still to come, buckets double up to ``TFC_BUCKET_MB``."""
        cap_max = max(self.bucket_elems // itemsize, 1)
        tail = max(int(float(os.environ.get("TFC_TAIL_BUCKET_MB", "4")) * (1 << 20)) // itemsize, 1)
        cap = min(tail, cap_max) if self.world > 1 else cap_max
        out, hi = [], len(sizes)
        while hi > 0:
            lo, acc = hi, 0
            while lo > 0 and acc < cap:
                lo -= 1
                acc += sizes[lo]
            out.append((lo, hi))
            hi = lo
            cap = min(cap * 2, cap_max)
        out.reverse()
        return out

    def _pick_algo(self, g: _Group, nbytes: int) -> int:
        """Schedule of one bucket by its size: flag-in-payload pushes while latency dominates
        (one-shot up to ``TFC_LL1_MAX_KB``, two-shot up to ``TFC_LL2_MAX_KB``), the in-switch
        reduction (or peer-load two-shot) once bandwidth does."""
        name = self.algo_name
        if self.world == 1:
            return comm_mod.ALGO_TWOSHOT            # "my shard" is everything: no peers, no barrier
        has_mc = bool(g.bufs and g.bufs[0].mc_ptr and g.bufs[1].mc_ptr)
        # two ranks: peer loads beat the switch (measured: 1 GiB 639 vs 430 GB/s busBW); four and more: NVLS
        big = comm_mod.ALGO_NVLS if (has_mc and self.world > 2) else comm_mod.ALGO_TWOSHOT
        cap1 = self.comm.ll_capacity(comm_mod.ALGO_LL1)
        cap2 = self.comm.ll_capacity(comm_mod.ALGO_LL2)
        if name == "auto":
            if nbytes <= min(cap1, int(os.environ.get("TFC_LL1_MAX_KB", "64")) << 10):
                return comm_mod.ALGO_LL1
            if nbytes <= min(cap2, int(os.environ.get("TFC_LL2_MAX_KB", "2048")) << 10):
                return comm_mod.ALGO_LL2
            if nbytes <= (1 << 20) and not cap1:
                return comm_mod.ALGO_ONESHOT
            return big
        algo = comm_mod.ALGO_NAMES[name]
        if algo == comm_mod.ALGO_NVLS and not has_mc:
            algo = comm_mod.ALGO_TWOSHOT
        if (algo == comm_mod.ALGO_LL1 and nbytes > cap1) or (algo == comm_mod.ALGO_LL2 and nbytes > cap2):
            algo = big
        return algo

    def _alloc_group(self, g: _Group, itemsize: int) -> None:
        n = g.total
        need_master = g.dtype != torch.float32
        if self.comm.uses_peer_memory:
            gb, pb = self.comm.alloc_pair(n * itemsize, n * itemsize)
            g.bufs = (gb, pb)
            g.grad_arena = gb.view(g.dtype, n)
            g.param_arena = pb.view(g.dtype, n)
            g.ctx = self.comm.make_ctx(gb, pb)
            for b in g.buckets:
                b.algo = self._pick_algo(g, b.length * itemsize)
                b.sharded = b.algo in comm_mod.SHARDED_ALGOS and self.world > 1
        else:
            g.grad_arena = self.comm.alloc_flat(n, g.dtype)        # in the shared segment on CPU ranks
            g.param_arena = torch.zeros(n, dtype=g.dtype, device=self.device)
        n_state = 0
        for b in g.buckets:
            b.s_off = n_state
            b.s_len = b.length // self.world if b.sharded else b.length
            n_state += b.s_len
        g.bucket_algos = [b.algo for b in g.buckets]
        g.algo = g.buckets[0].algo if g.buckets else 0
        g.sharded = any(b.sharded for b in g.buckets)
        code = self.opt.code
        z = lambda: torch.zeros(n_state, dtype=torch.float32, device=self.device)  # noqa: E731
        g.master = z() if need_master else None
        g.master_buf = None
        if need_master and g.sharded:
            # the fp32 master shards live in peer-mapped memory: a checkpoint written by the chief
            # alone (Keras' ModelCheckpoint) reads every rank's shard over NVLink, no collective
            from . import symm
            g.master_buf = symm.SymmetricBuffer(n_state * 4, self.device, self.comm.xchg, want_multicast=False)
            g.master = g.master_buf.view(torch.float32, n_state)
        g.s1 = z() if code != 0 else None
        g.s2 = z() if code in (1, 2, 3, 5) else None

    def _adopt_params(self, g: _Group) -> None:
        """Move every parameter into the arena and point .data / .grad at arena views."""
        with torch.no_grad():
            for (name, p), off in zip(g.params, g.offsets):
                n = p.numel()
                view = g.param_arena[off:off + n].view(p.shape)
                init = p.detach().to(device=self.device, dtype=torch.float32)
                if p.dim() == 4 and self.is_cuda:
                    # conv kernels live channels_last inside the arena (what cuDNN NHWC wants)
                    view = g.param_arena[off:off + n].view(
                        p.shape[0], p.shape[2], p.shape[3], p.shape[1]).permute(0, 3, 1, 2)
                view.copy_(init.to(g.dtype))
                p.data = view
                gview = g.grad_arena[off:off + n]
                if p.dim() == 4 and self.is_cuda:
                    gview = gview.view(p.shape[0], p.shape[2], p.shape[3], p.shape[1]).permute(0, 3, 1, 2)
                else:
                    gview = gview.view(p.shape)
                p.grad = gview
        self._refresh_master(g)

    def _state_range(self, g: _Group, b: _Bucket):
        """(arena offset, length, state offset) of the part of bucket b this rank updates."""
        if b.sharded:
            shard = b.length // self.world
            return b.off + self.rank * shard, shard, b.s_off
        return b.off, b.length, b.s_off

    def _pieces(self, b: _Bucket):
        """[(rank or None, arena lo, arena hi, state lo)]: whose state arrays hold which part of
        bucket b (None: every rank holds all of it)."""
        if not b.sharded:
            return [(None, b.off, b.off + b.length, b.s_off)]
        shard = b.length // self.world
        return [(r, b.off + r * shard, b.off + (r + 1) * shard, b.s_off) for r in range(self.world)]

    def _refresh_master(self, g: _Group) -> None:
        if g.master is None:
            return
        with torch.no_grad():
            for b in g.buckets:
                a_off, n, s_off = self._state_range(g, b)
                g.master[s_off:s_off + n].copy_(g.param_arena[a_off:a_off + n].float())

    def sync_from_root(self) -> None:
        """Initial-variable broadcast: replica 0's values everywhere (K10), then refresh masters."""
        if self.world > 1:
            for g in self.groups:
                if g.ctx is not None:
                    self.comm.broadcast_arena(g.ctx, g.dtype, 0, g.total, root=0)
                else:
                    self.comm.broadcast(g.param_arena, root=0)
            if self.is_cuda:
                torch.cuda.synchronize(self.device)
        for g in self.groups:
            self._refresh_master(g)

    # ------------------------------------------------------------------------------------------
    def _install_hooks(self) -> None:
        if self.is_cuda:
            from ..ops import bn
            bn.arm_scratch(self.device)
        for gi, g in enumerate(self.groups):
            bi, seen = 0, 0
            for (_name, p) in g.params:
                hook = self._make_hook(gi, bi)
                handle = p.register_post_accumulate_grad_hook(hook)
                self._hook_handles.append(handle)
                # kernels that own a parameter's gradient (fused batch-norm) may write it straight
                # into the arena view `p.grad` and call this instead of returning a gradient
                p._cb2_direct_grad = self.is_cuda
                p._cb2_grad_written = hook
                p._cb2_fp32 = self._fp32_accessor(gi, _name)
                seen += 1
                if seen == g.buckets[bi].n_params:
                    bi, seen = bi + 1, 0

    def _make_hook(self, gi: int, bi: int):
        def hook(_p):
            if not self._in_step:
                return
            # once per parameter and step: a kernel that wrote the gradient itself reports it here,
            # and autograd's post-accumulate hook may still fire for the same parameter afterwards
            key = id(_p)
            if key in self._written:
                return
            self._written.add(key)
            b = self.groups[gi].buckets[bi]
            b.ready += 1
            if b.ready == b.n_params and not b.launched:
                self._launch_bucket(gi, bi)
        return hook

    def remove_hooks(self) -> None:
        for h in self._hook_handles:
            h.remove()
        self._hook_handles = []
        for g in self.groups:
            for _n, p in g.params:
                for attr in ("_cb2_fp32", "_cb2_grad_written", "_cb2_direct_grad"):
                    if hasattr(p, attr):
                        try:
                            delattr(p, attr)
                        except AttributeError:
                            pass

    # ------------------------------------------------------------------------------------------
    # fp32 view of a parameter: checkpoints must hold the master values, not their bf16 rounding
    # ------------------------------------------------------------------------------------------
    def _fp32_accessor(self, gi: int, name: str):
        ref = weakref.ref(self)

        class _Acc:
            def get(_self):
                eng = ref()
                return None if eng is None else eng.master_value(gi, name)

            def set(_self, value):
                eng = ref()
                if eng is not None:
                    eng.set_master_value(gi, name, value)
        return _Acc()

    def _param_slot(self, gi: int, name: str):
        g = self.groups[gi]
        for (n, p), off in zip(g.params, g.offsets):
            if n == name:
                return g, p, off
        raise KeyError(name)

    def _to_logical(self, p, flat: torch.Tensor) -> torch.Tensor:
        """Arena order -> the parameter's logical shape (conv kernels are O,H,W,I in a CUDA arena)."""
        if p.dim() == 4 and self.is_cuda:
            return flat.view(p.shape[0], p.shape[2], p.shape[3], p.shape[1]).permute(0, 3, 1, 2)
        return flat.view(p.shape)

    def _to_arena_order(self, p, t: torch.Tensor) -> torch.Tensor:
        t = t.reshape(p.shape)
        if p.dim() == 4 and self.is_cuda:
            t = t.permute(0, 2, 3, 1)
        return t.contiguous().reshape(-1)

    def _gather_state(self, g: _Group, local: torch.Tensor, buf, lo: int, hi: int) -> torch.Tensor:
        """Elements [lo, hi) of a (possibly sharded) fp32 state vector, in arena order.  Sharded
        state that lives in peer-mapped memory (``buf``) is read from the peers directly."""
        from .symm import _wrap
        out = torch.empty(hi - lo, dtype=torch.float32, device=self.device)
        for b in g.buckets:
            for r, a_lo, a_hi, s_lo in self._pieces(b):
                a, z = max(lo, a_lo), min(hi, a_hi)
                if a >= z:
                    continue
                if r is None or r == self.rank:
                    src = local
                elif buf is None:
                    raise RuntimeError("sharded state without a peer mapping cannot be read by one rank")
                else:
                    src = _wrap(buf.peer_ptrs[r], buf.nbytes, self.device).view(torch.float32)
                k = s_lo + (a - a_lo)
                out[a - lo: z - lo].copy_(src[k: k + (z - a)])
        return out

    def master_value(self, gi: int, name: str) -> torch.Tensor:
        """The fp32 value of a parameter in its logical shape (CPU tensor): the master copy when the
        group has one, else the parameter itself.  No collective: safe to call on the chief alone."""
        g, p, off = self._param_slot(gi, name)
        n = p.numel()
        if g.master is None:
            return p.detach().float().cpu()
        if self.is_cuda:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        flat = self._gather_state(g, g.master, g.master_buf, off, off + n)
        return self._to_logical(p, flat).contiguous().cpu()

    def set_master_value(self, gi: int, name: str, value: torch.Tensor) -> None:
        """Overwrite a parameter with fp32 values: the master first (this rank's shard), then their
        rounding into the arena.  Every rank calls this with the same values (load_weights)."""
        g, p, off = self._param_slot(gi, name)
        n = p.numel()
        flat = self._to_arena_order(p, value.detach().to(self.device, torch.float32))
        with torch.no_grad():
            g.param_arena[off:off + n].copy_(flat.to(g.dtype))
            if g.master is None:
                return
            self._scatter_state(g, g.master, flat, off, n)
            self._master_exact.add((gi, name))

    def _scatter_state(self, g: _Group, dst: torch.Tensor, flat: torch.Tensor, off: int, n: int) -> None:
        """Write arena range [off, off + n) of a full-size vector into the parts of the state array
        ``dst`` this rank holds."""
        for b in g.buckets:
            for r, a_lo, a_hi, s_lo in self._pieces(b):
                if r is not None and r != self.rank:
                    continue
                a, z = max(a_lo, off), min(a_hi, off + n)
                if a < z:
                    dst[s_lo + (a - a_lo): s_lo + (z - a_lo)].copy_(flat[a - off: z - off])

    # ------------------------------------------------------------------------------------------
    def set_lr(self, lr: float) -> None:
        """Learning-rate schedules: write the new rate into the device hyper vector."""
        if float(lr) != self.opt.lr:
            self.opt.lr = float(lr)
            self.hyper[0:1].fill_(float(lr))

    def begin_step(self, n_aux: int = 0, direct: bool = True) -> None:
        """Call before backward: arms the bucket hooks and advances the step counter.  ``n_aux``:
        how many metric scalars of ``comm.aux`` the step wants summed over the replicas.
        ``direct``: kernels may write gradients straight into the arena and report them through
        ``_cb2_grad_written`` - true when the step ends in ``loss.backward()``; a GradientTape passes
        False, because ``torch.autograd.grad`` hands gradients back instead of accumulating them."""
        direct = bool(direct) and self.is_cuda
        if direct != self._direct:
            self._direct = direct
            for g in self.groups:
                for _n, p in g.params:
                    p._cb2_direct_grad = direct
        self._aux_pending = n_aux if (self.is_cuda and self.groups and self.groups[0].ctx is not None
                                      and os.environ.get("TFC_AUX_IN_BUCKET", "1") != "0") else 0
        self._aux_done = 0
        self._written = set()
        for g in self.groups:
            for b in g.buckets:
                b.ready, b.launched = 0, False
        if self._fault is not None and self._fault == (self.rank, self.steps):
            # fault injection for launcher / failure-detection tests: TFC_FAULT="<rank>:<step>"
            os._exit(17)
        optim_tick(self.hyper)
        if self.is_cuda:
            from ..ops import bn, elementwise
            elementwise.set_step_tensor(self.hyper[5:6])
            bn.reset_scratch(self.device)
        self._in_step = True

    def _launch_bucket(self, gi: int, bi: int) -> None:
        g = self.groups[gi]
        b = g.buckets[bi]
        b.launched = True
        if self.is_cuda:
            cur = torch.cuda.current_stream(self.device)
            self.comm_stream.wait_stream(cur)
            with torch.cuda.stream(self.comm_stream):
                self._bucket_update(g, b)
        else:
            self._bucket_update(g, b)

    def _bucket_update(self, g: _Group, b: _Bucket) -> None:
        code = self.opt.code
        if g.ctx is not None:
            _a, _n, s_off = self._state_range(g, b)
            # the step's metric scalars are final before backward starts: they ride in the first
            # bucket kernel that is launched (no launch of their own, no extra barrier rounds)
            n_aux, self._aux_pending = self._aux_pending, 0
            # device time of the bucket kernel for the metrics registry: a CUDA-event pair around it
            # whenever the step runs eagerly (the steps before the CUDA graph is captured, and every
            # step of a model that cannot be captured); inside a capture there is nothing to time
            timed = self._time_buckets and not torch.cuda.is_current_stream_capturing()
            if timed:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(self.comm_stream)
            self.comm.fused(g.ctx, algo=b.algo, opt=code, grad_dtype=g.dtype, param_dtype=g.dtype,
                            off=b.off, length=b.length, master=g.master, s1=g.s1, s2=g.s2,
                            state_off=s_off, hyper=self.hyper, zero_grads=True,
                            aux_len=n_aux, aux_out=self.aux_out if n_aux else None,
                            stream=self.comm_stream)
            if timed:
                e1.record(self.comm_stream)
                self._bucket_events.append((e0, e1))
                if len(self._bucket_events) > 4096:
                    del self._bucket_events[:2048]
            if n_aux:
                self._aux_done = n_aux
            return
        grad = g.grad_arena[b.off:b.off + b.length]
        if self.world > 1:
            t0 = time.perf_counter()
            self.comm.all_reduce_sum(grad)
            if not self.is_cuda:          # host-timed only where the call is synchronous
                self.allreduce_host_us.append((time.perf_counter() - t0) * 1e6)
        sl = slice(b.off, b.off + b.length)
        optim_flat(code, master=None if g.master is None else g.master[sl], grad=grad,
                   s1=None if g.s1 is None else g.s1[sl], s2=None if g.s2 is None else g.s2[sl],
                   param_out=g.param_arena[sl], hyper=self.hyper, zero_grad=True)

    def end_step(self, n_aux: int = 0) -> torch.Tensor | None:
        """Call after backward: flushes buckets whose hooks did not fire (unused parameters),
        reduces the first ``n_aux`` metric scalars of ``comm.aux`` across replicas and joins the
        side stream.  Returns the reduced scalars (device tensor) or None."""
        for gi, g in enumerate(self.groups):
            for bi, b in enumerate(g.buckets):
                if not b.launched:
                    self._launch_bucket(gi, bi)
        self._in_step = False
        out = None
        if n_aux:
            out = self.aux_out[:n_aux] if self._aux_done == n_aux else self._reduce_aux(n_aux)
        if self.is_cuda:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        self.steps += 1
        return out

    def _reduce_aux(self, n: int) -> torch.Tensor:
        g0 = self.groups[0] if self.groups else None
        if g0 is not None and g0.ctx is not None:
            cur = torch.cuda.current_stream(self.device)
            self.comm_stream.wait_stream(cur)
            with torch.cuda.stream(self.comm_stream):
                self.comm.fused(g0.ctx, algo=comm_mod.ALGO_ONESHOT, opt=0, grad_dtype=g0.dtype,
                                param_dtype=g0.dtype, off=0, length=0, master=None, s1=None,
                                s2=None, state_off=0, hyper=self.hyper, zero_grads=True,
                                aux_len=n, aux_out=self.aux_out, blocks=1,
                                stream=self.comm_stream)
            return self.aux_out[:n]
        if self.world > 1 and getattr(self.comm, "shm", None) is not None:
            self.aux_out[:n].copy_(self.comm.shm.reduce_aux(n))     # the scalars ride in the signal pad
            return self.aux_out[:n]
        self.aux_out[:n].copy_(self.comm.aux[:n])
        self.comm.aux[:n].zero_()
        if self.world > 1:
            self.comm.all_reduce_sum(self.aux_out[:n])
        return self.aux_out[:n]

    # ------------------------------------------------------------------------------------------
    def full_master_state(self) -> dict:
        """Gather the (possibly sharded) fp32 master + moments as full-size CPU tensors keyed by
        parameter name, each in the parameter's LOGICAL shape (a conv kernel is O,I,H,W whatever
        order the arena of this device keeps it in): what checkpoints store, so a run can resume
        under any strategy and on any device type."""
        import torch.distributed as dist
        out = {}
        for g in self.groups:
            full = {}
            for key, t in (("master", g.master), ("s1", g.s1), ("s2", g.s2)):
                if t is None:
                    continue
                parts = None
                if g.sharded:
                    parts = [torch.empty_like(t) for _ in range(self.world)]
                    dist.all_gather(parts, t.contiguous(), group=self.comm.group)
                buf = torch.empty(g.total, dtype=torch.float32, device=self.device)
                for b in g.buckets:
                    for r, a_lo, a_hi, s_lo in self._pieces(b):
                        src = t if r is None else parts[r]
                        buf[a_lo:a_hi] = src[s_lo:s_lo + (a_hi - a_lo)]
                full[key] = buf
            for (name, p), off in zip(g.params, g.offsets):
                n = p.numel()
                entry = {}
                for key, buf in full.items():
                    entry[key] = self._to_logical(p, buf[off:off + n].detach()).contiguous().cpu().clone()
                out[name] = entry
        out["__step__"] = float(self.hyper[5].item())
        out["__layout__"] = "logical"
        return out

    def load_master_state(self, state: dict) -> None:
        logical = state.get("__layout__") == "logical"
        for g in self.groups:
            for (name, p), off in zip(g.params, g.offsets):
                entry = state.get(name)
                if not entry:
                    continue
                n = p.numel()
                for key, t in (("master", g.master), ("s1", g.s1), ("s2", g.s2)):
                    if t is None or key not in entry:
                        continue
                    src = entry[key].to(self.device)
                    # files written before the layout tag hold the arena order of the device that
                    # wrote them (assumed to be this one's)
                    src = self._to_arena_order(p, src) if logical and src.dim() == p.dim() else src.reshape(-1)
                    self._scatter_state(g, t, src, off, n)
        if "__step__" in state:
            step = float(state["__step__"])
            self.hyper[5] = step
            self.hyper[6] = 1.0 - self.opt.beta1 ** step if step > 0 else 1.0
            self.hyper[7] = 1.0 - self.opt.beta2 ** step if step > 0 else 1.0

    def params_changed(self) -> None:
        """Call after parameters were overwritten from outside (load_weights/set_weights).  Masters
        that were just given exact fp32 values through ``set_master_value`` keep them; the others
        are re-derived from the arena."""
        exact = self._master_exact
        self._master_exact = set()
        keep = {}
        for gi, name in exact:
            keep[(gi, name)] = self.master_value(gi, name)
        for g in self.groups:
            self._refresh_master(g)
        for (gi, name), v in keep.items():
            self.set_master_value(gi, name, v)
        self._master_exact = set()

    def drain_allreduce_us(self) -> list:
        """Times (us) of the fused all-reduce + optimizer calls since the last drain: device times
        of the bucket kernels on GPU ranks (finished event pairs only - nothing is waited for),
        host times of the synchronous shared-memory / gloo calls on CPU ranks."""
        out, self.allreduce_host_us = self.allreduce_host_us, []
        keep = []
        for e0, e1 in self._bucket_events:
            if e1.query():
                out.append(e0.elapsed_time(e1) * 1e3)
            else:
                keep.append((e0, e1))
        self._bucket_events = keep
        return out

    def kernel_launches(self) -> int:
        return comm_mod.launches() + ops.optim.launches()
