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
            bucket_cap = max(self.bucket_elems // itemsize, 1)
            off, b_start, b_count = 0, 0, 0
            for _name, p in g.params:
                g.offsets.append(off)
                off += _round_up(p.numel(), 8)
                b_count += 1
                if off - b_start >= bucket_cap:
                    off = _round_up(off - b_start, align) + b_start
                    g.buckets.append(_Bucket(b_start, off - b_start, b_count))
                    b_start, b_count = off, 0
            if b_count:
                off = _round_up(off - b_start, align) + b_start
                g.buckets.append(_Bucket(b_start, off - b_start, b_count))
            g.total = off
            self._alloc_group(g, itemsize)
            self._adopt_params(g)
        self.sync_from_root()
        self._install_hooks()

    def _pick_algo(self, g: _Group, nbytes: int) -> int:
        name = self.algo_name
        if self.world == 1:
            return comm_mod.ALGO_TWOSHOT            # "my shard" is everything: no peers, no barrier
        has_mc = bool(g.bufs and g.bufs[0].mc_ptr and g.bufs[1].mc_ptr)
        if name == "auto":
            if nbytes <= (1 << 20):
                return comm_mod.ALGO_ONESHOT
            return comm_mod.ALGO_NVLS if has_mc else comm_mod.ALGO_TWOSHOT
        algo = comm_mod.ALGO_NAMES[name]
        if algo == comm_mod.ALGO_NVLS and not has_mc:
            algo = comm_mod.ALGO_TWOSHOT
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
            g.algo = self._pick_algo(g, n * itemsize)
            g.sharded = g.algo != comm_mod.ALGO_ONESHOT and self.world > 1
        else:
            g.grad_arena = self.comm.alloc_flat(n, g.dtype)        # in the shared segment on CPU ranks
            g.param_arena = torch.zeros(n, dtype=g.dtype, device=self.device)
        n_state = n // self.world if g.sharded else n
        code = self.opt.code
        z = lambda: torch.zeros(n_state, dtype=torch.float32, device=self.device)  # noqa: E731
        g.master = z() if need_master else None
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
        if g.sharded:
            shard = b.length // self.world
            return b.off + self.rank * shard, shard, b.off // self.world
        return b.off, b.length, b.off

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
                seen += 1
                if seen == g.buckets[bi].n_params:
                    bi, seen = bi + 1, 0

    def _make_hook(self, gi: int, bi: int):
        def hook(_p):
            if not self._in_step:
                return
            b = self.groups[gi].buckets[bi]
            b.ready += 1
            if b.ready == b.n_params and not b.launched:
                self._launch_bucket(gi, bi)
        return hook

    def remove_hooks(self) -> None:
        for h in self._hook_handles:
            h.remove()
        self._hook_handles = []

    # ------------------------------------------------------------------------------------------
    def set_lr(self, lr: float) -> None:
        """Learning-rate schedules: write the new rate into the device hyper vector."""
        if float(lr) != self.opt.lr:
            self.opt.lr = float(lr)
            self.hyper[0:1].fill_(float(lr))

    def begin_step(self) -> None:
        """Call before backward: arms the bucket hooks and advances the step counter."""
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
            self.comm.fused(g.ctx, algo=g.algo, opt=code, grad_dtype=g.dtype, param_dtype=g.dtype,
                            off=b.off, length=b.length, master=g.master, s1=g.s1, s2=g.s2,
                            state_off=s_off, hyper=self.hyper, zero_grads=True,
                            stream=self.comm_stream)
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
            out = self._reduce_aux(n_aux)
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
        parameter name: what checkpoints store, so a run can resume under any strategy."""
        import torch.distributed as dist
        out = {}
        for g in self.groups:
            full = {}
            for key, t in (("master", g.master), ("s1", g.s1), ("s2", g.s2)):
                if t is None:
                    continue
                if g.sharded:
                    parts = [torch.empty_like(t) for _ in range(self.world)]
                    dist.all_gather(parts, t, group=self.comm.group)
                    buf = torch.empty(g.total, dtype=torch.float32, device=self.device)
                    for b in g.buckets:
                        shard = b.length // self.world
                        for r in range(self.world):
                            s_off = b.off // self.world
                            buf[b.off + r * shard: b.off + (r + 1) * shard] = parts[r][s_off:s_off + shard]
                    full[key] = buf
                else:
                    full[key] = t
            for (name, p), off in zip(g.params, g.offsets):
                n = p.numel()
                entry = {}
                for key, buf in full.items():
                    entry[key] = buf[off:off + n].detach().cpu().clone()
                out[name] = entry
        out["__step__"] = float(self.hyper[5].item())
        return out

    def load_master_state(self, state: dict) -> None:
        for g in self.groups:
            for (name, p), off in zip(g.params, g.offsets):
                entry = state.get(name)
                if not entry:
                    continue
                n = p.numel()
                for key, t in (("master", g.master), ("s1", g.s1), ("s2", g.s2)):
                    if t is None or key not in entry:
                        continue
                    src = entry[key].to(self.device).reshape(-1)
                    if not g.sharded:
                        t[off:off + n].copy_(src)
                        continue
                    for b in g.buckets:
                        shard = b.length // self.world
                        lo = b.off + self.rank * shard
                        hi = lo + shard
                        a, z = max(lo, off), min(hi, off + n)
                        if a < z:
                            s_off = b.off // self.world
                            t[s_off + (a - lo): s_off + (z - lo)].copy_(src[a - off: z - off])
        if "__step__" in state:
            step = float(state["__step__"])
            self.hyper[5] = step
            self.hyper[6] = 1.0 - self.opt.beta1 ** step if step > 0 else 1.0
            self.hyper[7] = 1.0 - self.opt.beta2 ** step if step > 0 else 1.0

    def params_changed(self) -> None:
        """Call after parameters were overwritten from outside (load_weights/set_weights)."""
        for g in self.groups:
            self._refresh_master(g)

    def kernel_launches(self) -> int:
        return comm_mod.launches() + ops.optim.launches()
