"""``GradientTape`` for custom training loops (reference usage:
core/tests/testdata/mnist_example_using_ctl.py:103-131 - ``with tf.GradientTape() as tape:`` ...
``tape.gradient(loss, model.trainable_variables)`` ... ``optimizer.apply_gradients(zip(grads, vars))``).

torch records the graph whenever a parameter takes part in a computation, so the tape has nothing to
record.  What it does: it marks the start of a training step for the data-parallel engines of this
process (step counter for Adam's bias correction and the dropout stream, per-step scratch), and
``gradient`` differentiates with ``torch.autograd.grad``.  The update itself happens in
``Optimizer.apply_gradients`` (gradients are copied into the engine's flat arena, then one fused
all-reduce + optimizer kernel per bucket) or, overlapped with the backward pass, in
``Optimizer.minimize(loss, var_list, tape=tape)``.
"""
from __future__ import annotations

import torch


class GradientTape:
    def __init__(self, persistent=False, watch_accessed_variables=True):
        self.persistent = persistent
        self._used = False

    def __enter__(self):
        from ..parallel import engine as engine_mod
        for eng in engine_mod.live_engines():
            if not eng.in_step:
                eng.begin_step(direct=False)     # autograd.grad returns gradients: nothing may bypass it
        return self

    def __exit__(self, *exc):
        return False

    def watch(self, tensor):
        """Inputs are differentiable once they require grad; variables always are."""
        if isinstance(tensor, torch.Tensor) and tensor.is_floating_point() and not tensor.requires_grad:
            tensor.requires_grad_(True)
        return tensor

    def stop_recording(self):
        return torch.no_grad()

    def gradient(self, target, sources, output_gradients=None, unconnected_gradients="none"):
        """d target / d sources: a tensor for a tensor, a list for a list; ``None`` for a source
        the target does not depend on (zeros with ``unconnected_gradients="zero"``)."""
        if self._used and not self.persistent:
            raise RuntimeError("A non-persistent GradientTape can only be used to compute one set of "
                               "gradients (or jacobians)")
        self._used = True
        single = isinstance(sources, torch.Tensor)
        srcs = [sources] if single else list(sources)
        live = [s for s in srcs if s.requires_grad]
        got = torch.autograd.grad(target, live, grad_outputs=output_gradients, allow_unused=True,
                                  retain_graph=self.persistent) if live else ()
        by_id = {id(s): g for s, g in zip(live, got)}
        out = []
        for s in srcs:
            g = by_id.get(id(s))
            if g is None and unconnected_gradients == "zero":
                g = torch.zeros_like(s)
            out.append(g)
        return out[0] if single else out
