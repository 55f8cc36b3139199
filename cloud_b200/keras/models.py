"""Keras-shaped ``Model`` / ``Sequential`` on the DPEngine.

The reference's user scripts call ``model.compile(...)`` / ``model.fit(...)`` on TensorFlow and the
injected ``tf.distribute`` strategy does the rest (README.md:66-81; call stack SURVEY.md section
3.2).  This module is that "rest", B200-first:

* ``fit`` treats ``batch_size`` as the GLOBAL batch and splits it over replicas,
* per step: pinned-memory staging + async H2D on a copy stream, forward/backward in bf16 on the
  hand-written Dense / loss kernels (+ cuDNN for convolutions), gradient buckets reduced by the
  fused peer-memory all-reduce + optimizer kernel on a side stream, metrics summed across
  replicas by the same collective,
* the whole step is captured in a CUDA graph after a short eager warm-up (launch-bound MLP steps
  and batch-64 ResNet steps both gain from it),
* results are read back one step late so the host never stalls the device.
"""
from __future__ import annotations

import json
import os
import time

import numpy as np
import torch
import torch.nn as nn

from .. import ops
from ..utils.profiling import nvtx_range
from ..data import Dataset
from ..distribute import get_strategy
from ..distribute.strategy import DistributedDataset
from . import callbacks as cb_mod
from . import layers as L
from . import losses as losses_mod
from . import metrics as metrics_mod
from . import optimizers as opt_mod

AUX_LOSS, AUX_CORRECT, AUX_COUNT, AUX_METRIC0 = 0, 1, 2, 3


from .batch_sources import (_ArraySource, _grouped, _heartbeat, _IterSource, _NG, _np,  # noqa: F401
                            _Prefetcher, _SequenceAdapter)


class _Forked(list):
    """The two handles of a forked batch-norm output, handed out to its two consumers."""


_FORK_BN = os.environ.get("TFC_BN_FORK", "1") != "0"


class Model(nn.Module):
    """Functional / subclassable Keras model."""

    def __init__(self, inputs=None, outputs=None, name=None, **kwargs):
        super().__init__()
        self._weighted_targets = None
        self.name = name or L._auto_name(type(self).__name__)
        self.built = False
        self.stop_training = False
        self.optimizer = None
        self.loss = None
        self._metrics: list = []
        self._ratio_slots: dict = {}
        self._engine = None
        self._strategy = None
        self._graph = None
        self._graph_k = None          # graph of steps_per_execution steps
        self._graph_ok = True
        self.history = None
        self._nodes = None
        self._inputs = self._outputs = None
        self._layers = nn.ModuleList()
        self.run_eagerly = False
        if inputs is not None and outputs is not None:
            self._init_functional(inputs, outputs)

    # -- graph construction -------------------------------------------------------------------
    def _init_functional(self, inputs, outputs):
        self._inputs = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        self._outputs = list(outputs) if isinstance(outputs, (list, tuple)) else [outputs]
        order, seen = [], set()

        def visit(t):
            if id(t) in seen:
                return
            seen.add(id(t))
            for i in t.inputs:
                visit(i)
            order.append(t)
        for o in self._outputs:
            visit(o)
        self._nodes = order
        # fusion pass: an ImageNormalization whose only consumer is a ResNet-stem convolution
        # hands the raw bytes on (see ImageNormalization.defer_to_stem)
        consumers: dict = {}
        for t in order:
            for i in t.inputs:
                consumers.setdefault(id(i), []).append(t)
        outs = {id(o) for o in self._outputs}
        for t in order:
            if isinstance(t.layer, L.ImageNormalization) and id(t) not in outs:
                cons = consumers.get(id(t), [])
                if len(cons) == 1 and isinstance(cons[0].layer, L.Conv2D) and cons[0].layer.is_stem_geometry():
                    t.layer.defer_to_stem = True
        # a convolution whose output is read by anything but one BatchNormalization cannot hand its
        # operands on instead of its output (ops/convbn.py)
        for t in order:
            if isinstance(t.layer, L.Conv2D) and getattr(t.layer, "fold_candidate", False):
                cons = consumers.get(id(t), [])
                if id(t) in outs or len(cons) != 1 or not isinstance(cons[0].layer, L.BatchNormalization):
                    t.layer.fold_candidate = False
        # a batch-norm output with two consumers (the input of a residual block): in training its
        # two gradients are merged inside the batch-norm backward kernels (ops.bn_fork)
        self._fork_ids = {id(t) for t in order
                          if isinstance(t.layer, L.BatchNormalization) and id(t) not in outs
                          and len(consumers.get(id(t), [])) == 2}
        added = set()
        for t in order:
            if t.layer is not None and id(t.layer) not in added:
                added.add(id(t.layer))
                self._layers.append(t.layer)
        self.built = True

    @property
    def layers(self):
        return [l for l in self._layers if not isinstance(l, L.InputLayer)]

    def get_layer(self, name=None, index=None):
        if index is not None:
            return self.layers[index]
        for l in self._layers:
            if l.name == name:
                return l
        raise ValueError(f"No such layer: {name}.")

    @property
    def output(self):
        return self._outputs[0] if self._outputs and len(self._outputs) == 1 else self._outputs

    @property
    def input(self):
        return self._inputs[0] if self._inputs and len(self._inputs) == 1 else self._inputs

    @property
    def input_shape(self):
        return self._inputs[0].shape if self._inputs else None

    @property
    def output_shape(self):
        return self._outputs[0].shape if self._outputs else None

    @property
    def trainable_weights(self):
        return [p for p in self.parameters() if p.requires_grad]

    trainable_variables = trainable_weights

    @property
    def non_trainable_weights(self):
        """Frozen parameters and layer state that is not trained (batch-norm moving statistics)."""
        return [p for p in self.parameters() if not p.requires_grad] + list(self.buffers())

    non_trainable_variables = non_trainable_weights

    @property
    def weights(self):
        return self.trainable_weights + self.non_trainable_weights

    variables = weights

    def reset_metrics(self):
        for m in self._metrics:
            if hasattr(m, "reset_state"):
                m.reset_state()

    def reset_states(self):
        """Keras resets the state of stateful (recurrent) layers here; this front end has none."""

    @property
    def metrics_names(self):
        return ["loss"] + [m.name for m in self._metrics]

    # -- forward ------------------------------------------------------------------------------
    def call(self, x, training=False):
        if self._nodes is None:
            raise NotImplementedError("subclassed models must implement call()")
        xs = list(x) if isinstance(x, (list, tuple)) else [x]
        values = {id(t): v for t, v in zip(self._inputs, xs)}
        fork_ids = getattr(self, "_fork_ids", ()) if training and _FORK_BN else ()
        for t in self._nodes:
            if id(t) in values:
                continue
            ins = [v.pop() if isinstance(v, _Forked) else v for v in (values[id(i)] for i in t.inputs)]
            arg = ins if len(ins) > 1 else ins[0]
            if isinstance(t.layer, Model):
                values[id(t)] = t.layer.call(arg, training=training)
            else:
                values[id(t)] = t.layer.forward(arg, training=training)
            if id(t) in fork_ids and isinstance(values[id(t)], torch.Tensor):
                values[id(t)] = _Forked(ops.bn_fork(values[id(t)]))
        outs = [values[id(o)] for o in self._outputs]
        return outs[0] if len(outs) == 1 else outs

    def forward(self, x, training=False):
        return self.call(x, training=training)

    def _symbolic_call(self, inputs):
        ins = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        if not self.built:
            self._build_from_shape(ins[0].shape)
        return L.KerasTensor(self.compute_output_shape(ins[0].shape), layer=self, inputs=ins)

    def compute_output_shape(self, input_shape):
        return self._outputs[0].shape if self._outputs else input_shape

    def _build_from_shape(self, shape):
        pass

    def __call__(self, x, training=False, **kwargs):
        if isinstance(x, L.KerasTensor) or (isinstance(x, (list, tuple)) and x and
                                            isinstance(x[0], L.KerasTensor)):
            return self._symbolic_call(x)
        self._ensure_placed()
        if training and self._engine is None and not getattr(self, "_initial_sync_done", False):
            self._sync_initial_variables(x)
        xi = self._to_internal(x)
        was = self.training
        self.train(bool(training))
        try:
            out = self.call(xi, training=bool(training))
        finally:
            self.train(was)
        return _to_keras_layout(out)

    def _sync_initial_variables(self, sample_x):
        """Custom training loops call the model before any engine exists (the engine is built by the
        first ``apply_gradients``).  Mirrored variables start from replica 0's values
        (tf.distribute semantics), so the first training call broadcasts
        them; ``fit`` gets the same from the engine's own broadcast."""
        object.__setattr__(self, "_initial_sync_done", True)
        st = self.distribute_strategy
        if st.num_replicas_in_sync <= 1:
            return
        self._build_with_batch(sample_x)
        self._ensure_placed()
        with torch.no_grad():
            for t in list(self.parameters()) + list(self.buffers()):
                st.communicator.broadcast(t.data, root=0)

    # -- device / dtype plumbing --------------------------------------------------------------
    @property
    def distribute_strategy(self):
        if self._strategy is None:
            self._strategy = get_strategy()
        return self._strategy

    @property
    def _is_chief(self):
        return self.distribute_strategy.is_chief

    @property
    def device(self):
        return self.distribute_strategy.device

    def _ensure_placed(self):
        dev = self.device
        for t in list(self.parameters()) + list(self.buffers()):
            if t.device != dev:
                self.to(dev)
                break

    def _compute_dtype(self):
        if self._engine is not None:
            return self._engine.compute_dtype
        return torch.bfloat16 if self.device.type == "cuda" and _policy["name"] == "mixed_bfloat16" \
            else torch.float32

    def _to_internal(self, x, non_blocking=False):
        if isinstance(x, (list, tuple)):
            return [self._to_internal(t) for t in x]
        t = x if isinstance(x, torch.Tensor) else torch.as_tensor(_np(x))
        if t.device != self.device:
            t = t.to(self.device, non_blocking=non_blocking)
        if t.dim() == 4:
            t = t.permute(0, 3, 1, 2)        # NHWC batch -> NCHW-logical channels_last view
        if t.dtype == torch.uint8:
            return t                          # an ImageNormalization layer consumes raw bytes
        if not t.is_floating_point() and self._takes_token_ids():
            return t                          # ids for an Embedding must not be rounded to bf16
        if t.is_floating_point() or t.dtype in (torch.int32, torch.int64, torch.int16, torch.int8):
            t = t.to(self._compute_dtype())
        return t

    def _takes_token_ids(self) -> bool:
        flag = getattr(self, "_token_ids", None)
        if flag is None:
            flag = any(type(m).__name__ == "Embedding" for m in self.modules())
            object.__setattr__(self, "_token_ids", flag)
        return flag

    # -- compile ------------------------------------------------------------------------------
    def compile(self, optimizer="rmsprop", loss=None, metrics=None, run_eagerly=False,
                steps_per_execution=1, **kwargs):
        # Keras: "the number of batches to run during each call".  Here: that many training steps
        # are captured in ONE CUDA graph (inputs staged in a [K, batch, ...] buffer), which takes
        # the Python cost of a step (~0.25 ms) out of launch-bound models such as the README MLP.
        self.steps_per_execution = max(1, int(steps_per_execution or 1))
        self.optimizer = opt_mod.get(optimizer)
        self.loss = losses_mod.get(loss) if loss is not None else None
        self._metrics = [metrics_mod.get(m, self.loss) for m in (metrics or [])]
        # metrics that need more than one sum (precision: a denominator; AUC: counts per threshold)
        # keep their first value in their own slot and the rest behind all metrics, so the slots of the
        # common metrics never move
        self._ratio_slots, cursor = {}, AUX_METRIC0 + len(self._metrics)
        for i, m in enumerate(self._metrics):
            if m.extra_slots:
                self._ratio_slots[i] = (cursor, m.extra_slots)
                cursor += m.extra_slots
        from ..parallel.comm import AUX_FLOATS
        if cursor > AUX_FLOATS:
            raise ValueError("The metrics need {} scalars per step; {} fit in the step's scalar vector "
                             "(fewer AUC thresholds?)".format(cursor, AUX_FLOATS))
        self.run_eagerly = bool(run_eagerly)
        self._compiled_with = {"loss": loss, "metrics": list(metrics or [])}
        self._strategy = get_strategy()
        old = getattr(self, "_engine", None)
        if old is not None:
            # a re-compile (freeze / unfreeze, new optimizer) builds a new engine with new arenas:
            # the old one must stop reacting to gradients, and no CUDA graph recorded against its
            # arenas - single-step or steps_per_execution - may ever be replayed again
            old.remove_hooks()
            for g in old.groups:
                g.ctx = None
                g.bufs = ()
                g.grad_arena = g.param_arena = g.master = g.s1 = g.s2 = None
        self._engine = None
        self._graph = None
        self._graph_k = None
        self._regularized_layers = None
        self._weighted_targets = None

    def _build_with_batch(self, xb):
        """Run one forward on a single sample batch to create lazily-built weights."""
        if all(getattr(l, "built", True) for l in self._layers) and self.built:
            return
        with torch.no_grad():
            self.call(self._to_internal(xb[:1] if not isinstance(xb, (list, tuple)) else xb),
                      training=False)
        self.built = True

    def _ensure_engine(self, sample_x=None):
        if self._engine is not None:
            return self._engine
        if self.optimizer is None:
            raise RuntimeError("You must compile your model before training/testing. "
                               "Use `model.compile(optimizer, loss)`.")
        if sample_x is not None:
            self._build_with_batch(sample_x)
        self._ensure_placed()
        from ..parallel.engine import DPEngine
        st = self.distribute_strategy
        named = [(n, p) for n, p in self.named_parameters()]
        self._engine = DPEngine(named, self.optimizer.spec(), st.communicator,
                                compute_dtype=self._compute_dtype())
        self.optimizer._engine = self._engine         # custom loops on a compiled model share it
        self.optimizer._engine_vars = tuple(id(p) for _n, p in named if p.requires_grad)
        self._graph_ok = all(getattr(l, "graph_safe", True) for l in self.modules())
        cd = self._engine.compute_dtype
        if cd != torch.float32:
            # frozen matrices / conv kernels are not in the engine's arenas: give them the compute
            # dtype (and the channels_last layout the trainable ones have) here
            with torch.no_grad():
                for p in self.parameters():
                    if not p.requires_grad and p.dim() >= 2 and p.dtype != cd:
                        fmt = torch.channels_last if p.dim() == 4 else torch.preserve_format
                        p.data = p.data.to(dtype=cd, memory_format=fmt)
        self._mark_fusions()
        pending = getattr(self, "_pending_optimizer_state", None)
        if pending:
            self._engine.load_master_state(pending)
            self._pending_optimizer_state = None
        return self._engine

    def _mark_fusions(self):
        """Cross-layer fusions decided once the layers exist: Dense(relu) -> Dropout pairs (the
        dropout moves into the Dense op), and a softmax output layer whose probabilities nobody reads
        during fit() (fused cross-entropy + metrics computed by the same kernel)."""
        chain = []        # (layer, its only consumer layer or None)
        if getattr(self, "_seq", None):
            seq = [l for l in self._seq if not isinstance(l, L.InputLayer)]
            chain = list(zip(seq, seq[1:] + [None]))
            last = seq[-1] if seq else None
        elif self._nodes:
            consumers: dict = {}
            for t in self._nodes:
                for i in t.inputs:
                    consumers.setdefault(id(i), []).append(t)
            outs = {id(o) for o in self._outputs}
            for t in self._nodes:
                cons = consumers.get(id(t), [])
                nxt = cons[0].layer if (len(cons) == 1 and id(t) not in outs and len(cons[0].inputs) == 1) else None
                chain.append((t.layer, nxt))
            last = self._outputs[0].layer if len(self._outputs) == 1 else None
        else:
            last = None
        uses: dict = {}
        for layer, _n in chain:
            uses[id(layer)] = uses.get(id(layer), 0) + 1
        for layer, nxt in chain:
            if isinstance(layer, L.Dense):
                layer._fused_dropout = nxt if (isinstance(nxt, L.Dropout) and layer.activation_name == "relu"
                                               and uses.get(id(layer)) == 1 and uses.get(id(nxt)) == 1) else None
        ok = (isinstance(last, L.Dense) and last.activation_name == "softmax"
              and isinstance(self.loss, losses_mod.SparseCategoricalCrossentropy) and not self.loss.from_logits
              and all(m.from_fused_xent for m in self._metrics) and self._weighted_targets is None
              and self.device.type == "cuda")
        if os.environ.get("TFC_LOGITS_ONLY", "1") == "0":
            ok = False
        object.__setattr__(self, "_logits_only_layer", last if ok else None)

    # -- loss / metrics -----------------------------------------------------------------------
    def _loss_and_metrics(self, out, y, global_batch, aux):
        inv = 1.0 / float(global_batch)
        fused = False
        weights = None
        if self._weighted_targets is not None and self.training:
            # fit(sample_weight= / class_weight=): the weights ride in the last column of y
            weights = y[..., -1].float()
            y = y[..., :-1]
            if self._weighted_targets == "1d":
                y = y.squeeze(-1)
        if weights is None and isinstance(self.loss, losses_mod.SparseCategoricalCrossentropy) and out.dim() == 2:
            logits = self.loss.logits(out)
            if logits is not None:
                loss = ops.softmax_cross_entropy(logits, y.reshape(-1), grad_scale=inv, stats=aux, count=True,
                                                 unit_grad=True)
                fused = True
        if not fused:
            per = self.loss.per_sample(y, out.float() if out.dtype != torch.float32 else out)
            if weights is not None:
                per = per * weights             # Keras: sum(w * loss) / batch size
            s = per.sum()
            with torch.no_grad():
                aux[AUX_LOSS] += s.detach()
            loss = s * inv
        penalty = self._regularization_loss()
        if penalty is not None:
            # Keras adds the weight penalties to the loss it minimises and reports.  Gradients are
            # summed over replicas, so each adds its share (local batch / global batch) of one penalty.
            n = float(out.shape[0])
            with torch.no_grad():
                aux[AUX_LOSS] += penalty.detach() * n
            loss = loss + penalty * (n * inv)
        with torch.no_grad():
            if not fused:                     # the fused cross-entropy kernel counts its rows itself
                aux[AUX_COUNT] += float(out.shape[0])
            for i, m in enumerate(self._metrics):
                if fused and m.from_fused_xent:
                    continue
                if m.extra_slots:           # first value in the metric's slot, the others behind them all
                    vec = m.batch_vector(y, out)
                    start, n = self._ratio_slots[i]
                    aux[AUX_METRIC0 + i] += vec[0]
                    aux[start:start + n] += vec[1:]
                else:
                    aux[AUX_METRIC0 + i] += m.batch_sum(y, out)
        return loss, fused

    @property
    def trainable(self):
        return getattr(self, "_trainable", True)

    @trainable.setter
    def trainable(self, v):
        """``model.trainable = False`` freezes every layer (the reference's transfer-learning
        examples freeze the ResNet-50 base: call_run_within_script_with_keras_fit.py:79-87,
        dogs_classification.ipynb).  As in Keras, it takes effect at the next ``compile()``."""
        object.__setattr__(self, "_trainable", bool(v))
        for m in self.modules():
            if m is not self and isinstance(m, L.Layer):
                m.trainable = bool(v)

    @property
    def losses(self):
        """``tf.keras.Model.losses``: the weight penalties of all layers (see ``regularizers``)."""
        out = []
        for m in self.modules():
            if m is not self and isinstance(m, L.Layer) and not isinstance(m, Model):
                out.extend(m.losses)
        return out

    def _regularization_loss(self):
        layers = getattr(self, "_regularized_layers", None)
        if layers is None:              # found once per compile(): this runs every training step
            layers = self._regularized_layers = [
                m for m in self.modules() if m is not self and isinstance(m, L.Layer)
                and not isinstance(m, Model)
                and (m.kernel_regularizer is not None or m.bias_regularizer is not None)]
        if not layers:
            return None
        terms = [t for m in layers for t in m.losses]
        if not terms:
            return None
        total = terms[0]
        for t in terms[1:]:
            total = total + t
        return total

    def _aux_len(self):
        return AUX_METRIC0 + len(self._metrics) + sum(n for _s, n in self._ratio_slots.values())

    def _logs_from_aux(self, acc, fused, prefix=""):
        cnt = max(float(acc[AUX_COUNT]), 1.0)
        logs = {prefix + "loss": float(acc[AUX_LOSS]) / cnt}
        for i, m in enumerate(self._metrics):
            v = acc[AUX_CORRECT] if (fused and m.from_fused_xent) else acc[AUX_METRIC0 + i]
            if m.extra_slots:
                start, n = self._ratio_slots[i]
                logs[prefix + m.name] = m.reduce_vector([float(v)] + [float(t) for t in acc[start:start + n]])
            else:
                logs[prefix + m.name] = float(v) / cnt
        return logs

    # -- one training step --------------------------------------------------------------------
    def _train_step(self, xb, yb, global_batch=None):
        eng = self._engine
        eng.begin_step(self._aux_len())
        lo = getattr(self, "_logits_only_layer", None) if self._weighted_targets is None else None
        with nvtx_range("forward"):
            if lo is not None:
                L.LOGITS_ONLY["on"], L.LOGITS_ONLY["layer"] = True, lo
            try:
                out = self.call(xb, training=True)
            finally:
                L.LOGITS_ONLY["on"], L.LOGITS_ONLY["layer"] = False, None
        if global_batch is None:
            global_batch = out.shape[0] * eng.world
        with nvtx_range("loss"):
            loss, fused = self._loss_and_metrics(out, yb, global_batch, eng.comm.aux)
        self._last_fused = fused
        with nvtx_range("backward+bucketed all-reduce/optimizer"):
            loss.backward()
        with nvtx_range("end_step"):
            return eng.end_step(self._aux_len())

    def _graph_enabled(self):
        if self.run_eagerly or not self._graph_ok or self.device.type != "cuda":
            return False
        return os.environ.get("TFC_CUDA_GRAPH", "1") != "0"

    def _run_step(self, xb, yb, global_batch=None):
        """Eager for the first steps (and for odd shapes), then CUDA-graph replay."""
        _heartbeat()
        if not self._graph_enabled():
            return self._train_step(xb, yb, global_batch)
        g = self._graph
        if global_batch is None:        # same key whether or not the caller spelled it out
            global_batch = int(xb.shape[0]) * self._engine.world
        key = (tuple(xb.shape), xb.dtype, tuple(yb.shape), yb.dtype, int(global_batch))
        if g is None or g["key"] != key:
            if g is not None and g.get("frozen"):
                return self._train_step(xb, yb, global_batch)   # e.g. the short last batch
            self._graph = g = {"key": key, "warm": 0, "graph": None}
        if g["graph"] is None:
            g["warm"] += 1
            if g["warm"] <= int(os.environ.get("TFC_GRAPH_WARMUP", "3")):
                return self._train_step(xb, yb, global_batch)
            g["x"] = xb.clone()
            g["y"] = yb.clone()
            launches0 = ops.kernel_launches()
            torch.cuda.synchronize(self.device)
            graph = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                    g["out"] = self._train_step(g["x"], g["y"], global_batch)
            except Exception as e:  # capture refused: stay eager, say so once
                import warnings
                warnings.warn(f"cloud_b200: CUDA graph capture failed ({e}); running eagerly")
                self._graph_ok = False
                self._graph = None
                torch.cuda.synchronize(self.device)
                return self._train_step(xb, yb, global_batch)
            g["graph"] = graph
            g["frozen"] = True
            g["launches_per_step"] = ops.kernel_launches() - launches0
            # the capture itself did not execute: fall through to a replay for this batch
        g["x"].copy_(xb, non_blocking=True)
        g["y"].copy_(yb, non_blocking=True)
        g["graph"].replay()
        self._graph_replays = getattr(self, "_graph_replays", 0) + 1
        return g["out"]

    def _run_steps(self, xs, ys, global_batch, k):
        """``k`` training steps on the stacked batches ``xs[k, ...]`` / ``ys[k, ...]`` as one CUDA
        graph (``steps_per_execution``); returns the metric scalars summed over the k steps."""
        _heartbeat()
        n_aux = self._aux_len()
        if not self._graph_enabled():
            total = None
            for i in range(k):
                red = self._train_step(self._to_internal(xs[i]), ys[i], global_batch)[:n_aux].clone()
                total = red if total is None else total + red
            return total
        g = self._graph_k
        key = (tuple(xs.shape), xs.dtype, tuple(ys.shape), ys.dtype, int(global_batch), k)
        if g is None or g["key"] != key:
            self._graph_k = g = {"key": key, "warm": 0, "graph": None}
        if g["graph"] is None:
            g["warm"] += 1
            if g["warm"] <= 1 and (self._graph is None or self._graph.get("graph") is None):
                total = None                      # one eager pass (lazy builds, allocator warm-up)
                for i in range(k):
                    red = self._train_step(self._to_internal(xs[i]), ys[i], global_batch)[:n_aux].clone()
                    total = red if total is None else total + red
                return total
            g["x"], g["y"] = xs.clone(), ys.clone()
            g["sum"] = torch.zeros(n_aux, dtype=torch.float32, device=self.device)
            torch.cuda.synchronize(self.device)
            graph = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                    g["sum"].zero_()
                    for i in range(k):
                        red = self._train_step(self._to_internal(g["x"][i]), g["y"][i], global_batch)
                        g["sum"] += red[:n_aux]
            except Exception as e:
                import warnings
                warnings.warn(f"cloud_b200: CUDA graph capture of {k} steps failed ({e}); running eagerly")
                self._graph_ok = False
                self._graph_k = None
                torch.cuda.synchronize(self.device)
                return self._run_steps(xs, ys, global_batch, k)
            g["graph"] = graph
        g["x"].copy_(xs, non_blocking=True)
        g["y"].copy_(ys, non_blocking=True)
        g["graph"].replay()
        self._graph_replays = getattr(self, "_graph_replays", 0) + 1
        return g["sum"]

    # -- fit ----------------------------------------------------------------------------------
    def _make_source(self, x, y, batch_size, shuffle, steps, training=True):
        st = self.distribute_strategy
        world, rank = st.num_replicas_in_sync, st.replica_id
        if y is None and hasattr(x, "__getitem__") and hasattr(x, "__len__") and not isinstance(
                x, (np.ndarray, torch.Tensor, list, tuple, dict, Dataset, DistributedDataset)):
            x = _SequenceAdapter(x)                      # a keras.utils.Sequence (or look-alike)
        if isinstance(x, (Dataset, DistributedDataset)) or (
                not isinstance(x, (np.ndarray, torch.Tensor, list, tuple)) and hasattr(x, "__iter__")):
            if callable(getattr(x, "__next__", None)) and not hasattr(x, "__len__") \
                    and not isinstance(x, (Dataset, DistributedDataset)):
                pass
            return _IterSource(x, world, rank, steps)
        if isinstance(x, (list, tuple)) and not isinstance(x, np.ndarray):
            x = np.asarray(x)
        bs = batch_size or 32
        return _ArraySource(x, y, bs, shuffle and training, world, rank, steps=steps,
                            group=getattr(self, "steps_per_execution", 1) if training else 1)

    def fit(self, x=None, y=None, batch_size=None, epochs=1, verbose=1, callbacks=None,
            validation_split=0.0, validation_data=None, shuffle=True, initial_epoch=0,
            steps_per_epoch=None, validation_steps=None, validation_batch_size=None, **kwargs):
        if self.loss is None:
            raise RuntimeError("You must compile your model with a loss before calling fit().")
        if isinstance(x, Dataset) and kwargs.get("sample_weight") is None:
            # Keras: a dataset may yield (x, y, sample_weight) triples, and class_weight applies to the
            # labels of every batch it yields
            x, weighted = self._weighted_dataset(x, kwargs.get("class_weight"))
            weights = None
        else:
            weighted = None
            weights = self._sample_weights(x, y, kwargs.get("sample_weight"), kwargs.get("class_weight"))
        unknown = set(kwargs) - {"class_weight", "sample_weight", "workers", "use_multiprocessing",
                                 "max_queue_size", "validation_freq"}
        if unknown:
            raise TypeError("fit() got unexpected keyword arguments: {}".format(sorted(unknown)))
        if validation_split and validation_data is None and isinstance(x, (np.ndarray, torch.Tensor)):
            n_val = int(len(x) * validation_split)
            if n_val:
                validation_data = (x[-n_val:], y[-n_val:])
                x, y = x[:-n_val], y[:-n_val]
        if weights is not None:
            y_arr = _np(y).astype(np.float32)
            self._weighted_targets = "1d" if y_arr.ndim == 1 else "nd"
            y = np.concatenate([y_arr.reshape(len(y_arr), -1), weights.reshape(-1, 1)], axis=1)
        elif weighted is not None:
            self._weighted_targets = weighted
        try:
            return self._fit(x, y, batch_size, epochs, verbose, callbacks, validation_data, shuffle,
                             initial_epoch, steps_per_epoch, validation_steps, validation_batch_size, kwargs)
        finally:
            self._weighted_targets = None

    @staticmethod
    def _class_weight_table(class_weight, n_labels):
        table = np.ones(int(max(n_labels, max(class_weight) + 1)), dtype=np.float32)
        for cls, value in class_weight.items():
            table[int(cls)] = float(value)
        return table

    def _weighted_dataset(self, ds, class_weight):
        """A dataset whose elements are ``(x, y, sample_weight)`` and / or a ``class_weight`` table:
        returns the dataset mapped to ``(x, y with the weight as a last column)`` - the encoding
        ``_loss_and_metrics`` expects - and the target rank tag, or ``(ds, None)`` when there is
        nothing to weight."""
        first = next(iter(ds), None)
        triple = isinstance(first, tuple) and len(first) == 3
        if first is None or (not triple and class_weight is None):
            return ds, None
        if not isinstance(first, tuple) or len(first) < 2:
            raise ValueError("class_weight needs a dataset of (x, y) or (x, y, sample_weight) elements")
        y0 = _np(first[1])
        batched = bool(getattr(ds, "batched", True))
        label_rank = y0.ndim - (1 if batched else 0)
        tag = "1d" if label_rank == 0 else "nd"

        def fold(xb, yb, wb=None):
            ya = _np(yb).astype(np.float32)
            n = len(ya) if batched else 1
            w = np.ones(n, dtype=np.float32) if wb is None else _np(wb).astype(np.float32).reshape(-1)
            if class_weight is not None:
                labels = ya.reshape(n, -1)
                labels = labels.argmax(-1) if labels.shape[-1] > 1 else labels.reshape(-1)
                labels = labels.astype(np.int64)
                w = w * self._class_weight_table(class_weight, int(labels.max()) + 1)[labels]
            ext = np.concatenate([ya.reshape(n, -1), w.reshape(n, 1)], axis=1)
            return xb, (ext if batched else ext[0])

        return ds.map(fold), tag

    def _sample_weights(self, x, y, sample_weight, class_weight):
        """Per-sample loss weights of ``fit(sample_weight=..., class_weight=...)`` as one float32
        array (their product when both are given), or None.  Array inputs only."""
        if sample_weight is None and class_weight is None:
            return None
        if isinstance(x, (Dataset, DistributedDataset)) or y is None:
            raise NotImplementedError("sample_weight= needs array inputs (x, y): let the dataset yield "
                                      "(x, y, sample_weight) elements instead; class_weight works with "
                                      "a tfc.data.Dataset, not with a distributed dataset")
        n = len(y)
        w = np.ones(n, dtype=np.float32)
        if sample_weight is not None:
            sw = _np(sample_weight).astype(np.float32).reshape(-1)
            if len(sw) != n:
                raise ValueError("sample_weight has {} entries for {} samples".format(len(sw), n))
            w *= sw
        if class_weight is not None:
            labels = _np(y)
            if labels.ndim > 1 and labels.shape[-1] > 1:
                labels = labels.argmax(-1)              # one-hot targets
            labels = labels.reshape(-1).astype(np.int64)
            w *= self._class_weight_table(class_weight, int(labels.max()) + 1)[labels]
        return w

    def _fit(self, x, y, batch_size, epochs, verbose, callbacks, validation_data, shuffle, initial_epoch,
             steps_per_epoch, validation_steps, validation_batch_size, kwargs):
        source = self._make_source(x, y, batch_size, shuffle, steps_per_epoch)
        first = next(iter(_peek(source)))
        eng = self._ensure_engine(first[0])
        history = cb_mod.History()
        # Keras order: the user's callbacks, then the progress bar, then History - so the history
        # (and the printed line) include what callbacks add to the logs, e.g. "lr"
        cbs = list(callbacks or [])
        if verbose and self._is_chief:
            cbs.append(cb_mod.ProgbarLogger(verbose))
        cbs.append(history)
        try:
            n_steps = len(source)
        except TypeError:                  # unknown until the first epoch has been counted
            n_steps = None
        params = {"epochs": epochs, "steps": n_steps, "verbose": verbose}
        cbl = cb_mod.CallbackList(cbs, self, params)
        batch_logs = _needs_batch_logs(cbs)
        self.stop_training = False
        self.train(True)
        self._resume_from_epoch = 0
        cbl.call("on_train_begin", {})
        initial_epoch = max(initial_epoch, int(self._resume_from_epoch))      # set by BackupAndRestore
        n_aux = self._aux_len()
        self._h2d_bytes = 0
        self._d2h_bytes = 0
        logs = {}
        for epoch in range(initial_epoch, epochs):
            cbl.call("on_epoch_begin", epoch, {})
            eng.set_lr(self.optimizer.current_lr())
            acc = torch.zeros(n_aux, dtype=torch.float64, device=self.device)
            pending = None           # (step index, device tensor of reduced aux, event)
            pf = _Prefetcher(source, self.device)
            step = -1
            epoch_t0 = time.perf_counter()
            # where the host thread spends the epoch (bench.py reports it next to the e2e number)
            prof = self._fit_profile = {"wait_data_s": 0.0, "launch_s": 0.0, "read_back_s": 0.0, "steps": 0}
            t_mark = epoch_t0
            for group in _grouped(pf, getattr(self, "steps_per_execution", 1)):
                now = time.perf_counter()
                prof["wait_data_s"] += now - t_mark
                step += len(group)
                if batch_logs:
                    cbl.call("on_train_batch_begin", step, {})
                cur = torch.cuda.current_stream(self.device) if self.device.type == "cuda" else None
                for (xd, yd, ev, _ng) in group:
                    if ev is not None:
                        cur.wait_event(ev)
                        xd.record_stream(cur)          # allocated on the loader's copy stream
                        if yd is not None:
                            yd.record_stream(cur)
                n_global = group[0][3]
                pre_k = getattr(n_global, "k", 1)
                if pre_k > 1:                       # the source already stacked k batches
                    step += pre_k - 1
                    red = self._run_steps(group[0][0], group[0][1].to(self.device), int(n_global), pre_k)
                    self.optimizer.iterations += pre_k - 1
                elif len(group) == 1:
                    xd, yd = group[0][0], group[0][1]
                    xi = self._to_internal(xd)
                    yi = yd.to(self.device) if yd is not None else None
                    red = self._run_step(xi, yi, n_global)
                else:
                    xs = torch.stack([b[0].to(self.device) for b in group])
                    ys = torch.stack([b[1].to(self.device) for b in group])
                    red = self._run_steps(xs, ys, n_global, len(group))
                self.optimizer.iterations += len(group)
                if self.optimizer._schedule is not None:
                    eng.set_lr(self.optimizer.current_lr())
                acc += red[:n_aux]
                t_launched = time.perf_counter()
                prof["launch_s"] += t_launched - now
                if batch_logs:
                    # read step i-1 while step i runs: the host never drains the device
                    snap = _HostSnap(acc, self.device)
                    if pending is not None:
                        self._emit_batch(cbl, pending)
                    pending = (step, snap)
                t_mark = time.perf_counter()
                prof["read_back_s"] += t_mark - t_launched
                prof["steps"] = step + 1
                if self.stop_training:
                    break
            if pending is not None:
                self._emit_batch(cbl, pending)
            self._h2d_bytes += pf.h2d_bytes
            host = acc.cpu()
            self._d2h_bytes += host.numel() * 8
            self._record_epoch_metrics(eng, step + 1, time.perf_counter() - epoch_t0,
                                       float(host[AUX_COUNT]))
            logs = self._logs_from_aux(host, getattr(self, "_last_fused", False)) if step >= 0 else {}
            if validation_data is not None and (epoch + 1) % max(1, int(kwargs.get("validation_freq") or 1)) == 0:
                vx, vy = (validation_data[0], validation_data[1]) \
                    if isinstance(validation_data, (tuple, list)) else (validation_data, None)
                val = self.evaluate(vx, vy, batch_size=validation_batch_size or batch_size,
                                    steps=validation_steps, verbose=0, return_dict=True)
                logs.update({"val_" + k: v for k, v in val.items()})
                self.train(True)
            cbl.call("on_epoch_end", epoch, logs)
            if self.stop_training:
                break
        cbl.call("on_train_end", logs)
        self.train(False)
        self.history = history
        return history

    def _record_epoch_metrics(self, eng, steps, seconds, samples):
        """Feeds the C++ metrics registry (cloud_b200.monitoring); exported only if enabled."""
        if steps <= 0 or seconds <= 0:
            return
        try:
            from .. import monitoring as mon
            if not getattr(self, "_mon_ready", False):
                mon.define_defaults()
                mon.start()
                self._mon_ready = True
            mon.observe(mon.STEP_TIME, seconds / steps * 1e6)
            mon.set_gauge(mon.IMAGES_PER_SEC, int(samples / seconds))
            prof = getattr(self, "_fit_profile", None)
            if prof and prof.get("steps"):
                mon.observe(mon.DATA_WAIT, prof["wait_data_s"] / prof["steps"] * 1e6)
            for us in eng.drain_allreduce_us()[-4096:]:
                mon.observe(mon.ALLREDUCE_TIME, us)
            n_buckets = sum(len(g.buckets) for g in eng.groups)
            mon.increment(mon.BUCKET_LAUNCHES, steps * n_buckets)
            nbytes = sum(g.total * (2 if g.dtype == torch.bfloat16 else 4) for g in eng.groups)
            if eng.world > 1 or eng.is_cuda:          # on one GPU: the bytes the fused kernel updates
                mon.increment(mon.BYTES_REDUCED, steps * nbytes)
        except Exception:      # metrics must never break training
            pass

    def _emit_batch(self, cbl, pending):
        step, snap = pending
        host = snap.get()
        self._d2h_bytes += host.numel() * 8
        cbl.call("on_train_batch_end", step,
                 self._logs_from_aux(host, getattr(self, "_last_fused", False)))

    def train_on_batch(self, x, y=None, reset_metrics=True, return_dict=False, **kwargs):
        eng = self._ensure_engine(torch.as_tensor(_np(x)) if not isinstance(x, torch.Tensor) else x)
        self.train(True)
        eng.set_lr(self.optimizer.current_lr())
        xi = self._to_internal(x)
        yi = torch.as_tensor(_np(y)).to(self.device) if not isinstance(y, torch.Tensor) else y.to(self.device)
        red = self._run_step(xi, yi)
        self.optimizer.iterations += 1
        host = red[:self._aux_len()].double().cpu()
        logs = self._logs_from_aux(host, getattr(self, "_last_fused", False))
        self.train(False)
        if return_dict:
            return logs
        vals = list(logs.values())
        return vals[0] if len(vals) == 1 else vals

    # -- evaluate / predict -------------------------------------------------------------------
    def evaluate(self, x=None, y=None, batch_size=None, verbose=1, steps=None, callbacks=None,
                 return_dict=False, **kwargs):
        if self.loss is None:
            raise RuntimeError("You must compile your model before training/testing.")
        source = self._make_source(x, y, batch_size, False, steps, training=False)
        first = next(iter(_peek(source)))
        eng = self._ensure_engine(first[0])
        self.train(False)
        n_aux = self._aux_len()
        aux = torch.zeros(max(n_aux, 8), dtype=torch.float32, device=self.device)
        fused = False
        with torch.no_grad():
            for xd, yd, ev, _ng in _Prefetcher(source, self.device):
                if ev is not None:
                    torch.cuda.current_stream(self.device).wait_event(ev)
                out = self.call(self._to_internal(xd), training=False)
                _l, fused = self._loss_and_metrics(out, yd.to(self.device), 1, aux)
        if eng.world > 1:
            eng.comm.all_reduce_sum(aux)
        logs = self._logs_from_aux(aux.double().cpu(), fused)
        if verbose and self._is_chief:
            print(" - ".join(f"{k}: {v:.4f}" for k, v in logs.items()), flush=True)
        if return_dict:
            return logs
        vals = list(logs.values())
        return vals[0] if len(vals) == 1 else vals

    def test_on_batch(self, x, y=None, return_dict=False, **kwargs):
        return self.evaluate(x, y, batch_size=len(x), verbose=0, return_dict=return_dict)

    def predict(self, x, batch_size=None, verbose=0, steps=None, **kwargs):
        self._ensure_placed()
        self.train(False)
        outs = []
        if isinstance(x, (Dataset, DistributedDataset)):
            batches = (b[0] if isinstance(b, (tuple, list)) else b for b in x)
        else:
            arr = x if isinstance(x, torch.Tensor) else torch.as_tensor(_np(x))
            bs = batch_size or 32
            batches = (arr[i:i + bs] for i in range(0, len(arr), bs))
        with torch.no_grad():
            for i, b in enumerate(batches):
                if steps is not None and i >= steps:
                    break
                out = _to_keras_layout(self.call(self._to_internal(b), training=False))
                outs.append(out.float().cpu())
        return torch.cat(outs).numpy() if outs else np.zeros((0,))

    def predict_on_batch(self, x):
        return self.predict(x, batch_size=len(x))

    # -- weights ------------------------------------------------------------------------------
    def get_weights(self):
        out = []
        for l in self._weight_layers():
            out.extend(l.get_weights())
        return out

    def set_weights(self, weights):
        i = 0
        for l in self._weight_layers():
            n = len(l._keras_weight_specs)
            l.set_weights(weights[i:i + n])
            i += n
        if self._engine is not None:
            self._engine.params_changed()

    def _weight_layers(self):
        out = []
        for l in self._layers:
            if isinstance(l, Model):
                out.extend(l._weight_layers())
            elif getattr(l, "_keras_weight_specs", None):
                out.append(l)
        return out

    def count_params(self):
        return sum(l.count_params() for l in self._weight_layers())

    def save_weights(self, filepath, overwrite=True, save_format=None, **kwargs):
        from . import saving
        saving.save_weights(self, filepath)

    def load_weights(self, filepath, by_name=False, **kwargs):
        from . import saving
        saving.load_weights(self, filepath, by_name=by_name)
        return _LoadStatus()

    def save(self, filepath, overwrite=True, include_optimizer=True, save_format=None, **kwargs):
        from . import saving
        saving.save_model(self, filepath, include_optimizer=include_optimizer)

    # -- config -------------------------------------------------------------------------------
    def get_config(self):
        nodes = []
        names = {}
        for t in self._nodes or []:
            names[id(t)] = t.layer.name
            nodes.append({"class_name": type(t.layer).__name__, "name": t.layer.name,
                          "config": _layer_config(t.layer),
                          "inbound": [names[id(i)] for i in t.inputs],
                          "shape": list(t.shape)})
        return {"name": self.name, "nodes": nodes,
                "inputs": [names[id(t)] for t in self._inputs or []],
                "outputs": [names[id(t)] for t in self._outputs or []]}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        tensors = {}
        for node in config["nodes"]:
            if node["class_name"] == "InputLayer":
                shape = tuple(node["shape"][1:])
                tensors[node["name"]] = L.Input(shape=shape, name=node["name"])
                continue
            layer = _layer_from_config(node["class_name"], node["config"], custom_objects)
            ins = [tensors[n] for n in node["inbound"]]
            tensors[node["name"]] = layer(ins if len(ins) > 1 else ins[0])
        return cls(inputs=[tensors[n] for n in config["inputs"]],
                   outputs=[tensors[n] for n in config["outputs"]], name=config.get("name"))

    def to_json(self):
        return json.dumps({"class_name": type(self).__name__, "config": self.get_config()})

    def summary(self, print_fn=print):
        print_fn(f'Model: "{self.name}"')
        print_fn("_" * 65)
        print_fn(f"{'Layer (type)':<29}{'Output Shape':<26}{'Param #':<10}")
        print_fn("=" * 65)
        total = 0
        for l in self._layers:
            n = l.count_params() if hasattr(l, "count_params") else 0
            total += n
            shape = getattr(l, "output_shape_", None) or getattr(l, "output_shape", None)
            print_fn(f"{l.name + ' (' + type(l).__name__ + ')':<29}{str(shape):<26}{n:<10}")
        print_fn("=" * 65)
        print_fn(f"Total params: {total:,}")


class _LoadStatus:
    def expect_partial(self):
        return self

    def assert_consumed(self):
        return self

    def assert_existing_objects_matched(self):
        return self


class _HostSnap:
    """Async device->host copy of a small tensor into pinned memory, read later."""

    _ring: dict = {}       # (shape, dtype) -> [pinned buffers, events, cursor]; pinning is slow

    def __init__(self, t, device):
        if device.type == "cuda":
            key = (tuple(t.shape), t.dtype)
            ring = _HostSnap._ring.get(key)
            if ring is None:
                ring = [[torch.empty(t.shape, dtype=t.dtype).pin_memory() for _ in range(4)],
                        [torch.cuda.Event() for _ in range(4)], 0]
                _HostSnap._ring[key] = ring
            i = ring[2] % 4
            ring[2] += 1
            self.buf, self.ev = ring[0][i], ring[1][i]
            self.buf.copy_(t, non_blocking=True)
            self.ev.record()
        else:
            self.buf, self.ev = t.clone(), None

    def get(self):
        if self.ev is not None:
            self.ev.synchronize()
            return self.buf.clone()      # the ring slot is reused four snapshots later
        return self.buf


def _needs_batch_logs(cbs) -> bool:
    base = cb_mod.Callback
    for cb in cbs:
        if isinstance(cb, (cb_mod.History, cb_mod.ProgbarLogger)):
            continue
        for name in ("on_train_batch_end", "on_batch_end", "on_train_batch_begin", "on_batch_begin"):
            fn = getattr(cb, name)
            if getattr(fn, "__func__", None) is not getattr(base, name):
                return True
    return False


def _peek(source):
    """First batch of a source without disturbing array sources (datasets restart anyway)."""
    if isinstance(source, _ArraySource):
        yield source.x[:1], None
        return
    it = iter(source)
    first = next(it)
    source._it = None
    yield first


def _to_keras_layout(out):
    if isinstance(out, (list, tuple)):
        return [_to_keras_layout(o) for o in out]
    if isinstance(out, torch.Tensor) and out.dim() == 4:
        return out.permute(0, 2, 3, 1)
    return out


def _layer_config(layer):
    if isinstance(layer, Model):
        return {"model": layer.get_config(), "kind": type(layer).__name__}
    cfg = layer.get_config()
    if isinstance(layer, L.InputLayer):
        cfg["input_shape"] = list(layer.input_shape_[1:]) if layer.input_shape_ else None
    return cfg


def _layer_from_config(class_name, config, custom_objects=None):
    if class_name in ("Model", "Sequential", "Functional"):
        cls = Sequential if config.get("kind") == "Sequential" else Model
        return cls.from_config(config["model"], custom_objects)
    table = dict(L.LAYER_CLASSES)
    table.update(custom_objects or {})
    if class_name not in table:
        raise ValueError(f"Unknown layer: {class_name}")
    cfg = dict(config)
    return table[class_name](**cfg)


class Sequential(Model):
    def __init__(self, layers=None, name=None):
        super().__init__(name=name)
        self._seq: list = []
        for l in layers or []:
            self.add(l)

    def add(self, layer):
        if isinstance(layer, L.KerasTensor):       # Sequential([Input(...), ...])
            layer = layer.layer
        self._seq.append(layer)
        self._layers.append(layer)
        if len(self._seq) == 1 and getattr(layer, "_declared_input_shape", None):
            self._build_from_shape(layer._declared_input_shape)
        elif self.built and self._outputs:
            self._outputs = [layer(self._outputs[0])]
            self._nodes = None

    def _build_from_shape(self, shape):
        """Static shape inference through the stack (creates the weights, Keras-style)."""
        t = L.KerasTensor(tuple(shape))
        self._inputs = [t]
        for l in self._seq:
            if isinstance(l, L.InputLayer):
                continue
            t = l(t)
        self._outputs = [t]
        self.built = True

    def call(self, x, training=False):
        for l in self._seq:
            if isinstance(l, L.InputLayer):
                continue
            x = l.call(x, training=training) if isinstance(l, Model) else l.forward(x, training=training)
        if not self.built:
            self.built = True
        return x

    def _build_with_batch(self, xb):
        if self.built and all(getattr(l, "built", True) for l in self._seq):
            return
        t = xb if isinstance(xb, torch.Tensor) else torch.as_tensor(_np(xb))
        self._build_from_shape((None,) + tuple(t.shape[1:]))

    def build(self, input_shape=None):
        if input_shape is not None:
            self._build_from_shape(tuple(input_shape))

    def pop(self):
        l = self._seq.pop()
        self._layers = nn.ModuleList(self._seq)
        return l

    def get_config(self):
        return {"name": self.name, "kind": "Sequential",
                "input_shape": list(self._inputs[0].shape) if self._inputs else None,
                "layers": [{"class_name": type(l).__name__, "config": _layer_config(l)}
                           for l in self._seq if not isinstance(l, L.InputLayer)]}

    @classmethod
    def from_config(cls, config, custom_objects=None):
        m = cls(name=config.get("name"))
        for ld in config["layers"]:
            m.add(_layer_from_config(ld["class_name"], ld["config"], custom_objects))
        if config.get("input_shape"):
            m._build_from_shape(tuple(config["input_shape"]))
        return m


_policy = {"name": "mixed_bfloat16"}


class mixed_precision:
    """``tf.keras.mixed_precision``-like global policy: 'mixed_bfloat16' (default on B200: bf16
    compute, fp32 master weights) or 'float32'."""

    @staticmethod
    def set_global_policy(name):
        if name not in ("mixed_bfloat16", "float32"):
            raise ValueError("policy must be 'mixed_bfloat16' or 'float32'")
        _policy["name"] = name

    @staticmethod
    def global_policy():
        return _policy["name"]


def load_model(filepath, custom_objects=None, compile=True):
    from . import saving
    return saving.load_model(filepath, custom_objects=custom_objects, compile=compile)


def model_from_json(text, custom_objects=None):
    d = json.loads(text)
    cls = Sequential if d["config"].get("kind") == "Sequential" else Model
    return cls.from_config(d["config"], custom_objects)
