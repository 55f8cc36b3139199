"""Keras-shaped optimizers.  They only *describe* the update; the update itself is executed by
the DPEngine's fused kernels (``csrc/kernels/comm.cu`` / ``elementwise.cu``), one launch per
gradient bucket instead of one ``ResourceApplyAdam`` per variable.

Reference usage: ``Adam()`` README.md:78, ``Adam(lr=hp.get('learning_rate'))``
tuner/tests/integration/tuner_integration_test.py:75, ``SGD(0.002)`` cloud_fit client_test.py:87-89,
``RMSprop(learning_rate=ExponentialDecay(...))`` core/tests/testdata/keras_tuner_cifar_example.py:66-76.
"""
from __future__ import annotations

import math

from ..parallel.engine import OptimizerSpec


class LearningRateSchedule:
    def __call__(self, step):
        raise NotImplementedError

    def get_config(self):
        return {}


class ExponentialDecay(LearningRateSchedule):
    def __init__(self, initial_learning_rate, decay_steps, decay_rate, staircase=False, name=None):
        self.initial_learning_rate = float(initial_learning_rate)
        self.decay_steps = decay_steps
        self.decay_rate = decay_rate
        self.staircase = staircase

    def __call__(self, step):
        p = step / self.decay_steps
        if self.staircase:
            p = float(int(p))
        return self.initial_learning_rate * (self.decay_rate ** p)

    def get_config(self):
        return {"initial_learning_rate": self.initial_learning_rate, "decay_steps": self.decay_steps,
                "decay_rate": self.decay_rate, "staircase": self.staircase}


class PiecewiseConstantDecay(LearningRateSchedule):
    def __init__(self, boundaries, values, name=None):
        self.boundaries, self.values = list(boundaries), list(values)

    def __call__(self, step):
        for b, v in zip(self.boundaries, self.values):
            if step <= b:
                return v
        return self.values[-1]


    def get_config(self):
        return {"boundaries": self.boundaries, "values": self.values}


class CosineDecay(LearningRateSchedule):
    """``initial * ((1 - alpha) * 0.5 * (1 + cos(pi * min(step, decay_steps) / decay_steps)) + alpha)``,
    optionally after a linear warm-up from ``initial`` to ``warmup_target`` over ``warmup_steps``."""

    def __init__(self, initial_learning_rate, decay_steps, alpha=0.0, name=None, warmup_target=None,
                 warmup_steps=0):
        self.initial_learning_rate = float(initial_learning_rate)
        self.decay_steps, self.alpha = int(decay_steps), float(alpha)
        self.warmup_target, self.warmup_steps = warmup_target, int(warmup_steps)

    def __call__(self, step):
        peak = self.initial_learning_rate
        if self.warmup_target is not None and self.warmup_steps > 0:
            if step < self.warmup_steps:
                return peak + (self.warmup_target - peak) * step / self.warmup_steps
            peak, step = float(self.warmup_target), step - self.warmup_steps
        frac = min(step, self.decay_steps) / float(self.decay_steps)
        return peak * ((1.0 - self.alpha) * 0.5 * (1.0 + math.cos(math.pi * frac)) + self.alpha)

    def get_config(self):
        return {"initial_learning_rate": self.initial_learning_rate, "decay_steps": self.decay_steps,
                "alpha": self.alpha, "warmup_target": self.warmup_target, "warmup_steps": self.warmup_steps}


class PolynomialDecay(LearningRateSchedule):
    def __init__(self, initial_learning_rate, decay_steps, end_learning_rate=0.0001, power=1.0, cycle=False,
                 name=None):
        self.initial_learning_rate, self.end_learning_rate = float(initial_learning_rate), float(end_learning_rate)
        self.decay_steps, self.power, self.cycle = int(decay_steps), float(power), cycle

    def __call__(self, step):
        steps = self.decay_steps
        if self.cycle:
            steps = steps * max(1.0, math.ceil(step / steps))
        else:
            step = min(step, steps)
        return (self.initial_learning_rate - self.end_learning_rate) * (1 - step / steps) ** self.power \
            + self.end_learning_rate

    def get_config(self):
        return {"initial_learning_rate": self.initial_learning_rate, "decay_steps": self.decay_steps,
                "end_learning_rate": self.end_learning_rate, "power": self.power, "cycle": self.cycle}


class InverseTimeDecay(LearningRateSchedule):
    def __init__(self, initial_learning_rate, decay_steps, decay_rate, staircase=False, name=None):
        self.initial_learning_rate = float(initial_learning_rate)
        self.decay_steps, self.decay_rate, self.staircase = decay_steps, decay_rate, staircase

    def __call__(self, step):
        p = step / self.decay_steps
        if self.staircase:
            p = float(int(p))
        return self.initial_learning_rate / (1.0 + self.decay_rate * p)

    def get_config(self):
        return {"initial_learning_rate": self.initial_learning_rate, "decay_steps": self.decay_steps,
                "decay_rate": self.decay_rate, "staircase": self.staircase}


class schedules:  # tf.keras.optimizers.schedules.*
    LearningRateSchedule = LearningRateSchedule
    ExponentialDecay = ExponentialDecay
    PiecewiseConstantDecay = PiecewiseConstantDecay
    CosineDecay = CosineDecay
    PolynomialDecay = PolynomialDecay
    InverseTimeDecay = InverseTimeDecay


class _HyperParameter(float):
    """What ``optimizer.lr`` / ``optimizer.learning_rate`` read as: a float that also answers the
    ``tf.Variable`` calls user callbacks make on it (``.numpy()``, ``.assign(v)``) - e.g. the
    ``PrintLR`` callback of the reference's examples (testdata/mnist_example_using_fit.py:82-88)."""

    def __new__(cls, value, owner=None):
        self = super().__new__(cls, value)
        self._owner = owner
        return self

    def numpy(self):
        return float(self)

    def assign(self, value):
        if self._owner is not None:
            self._owner.learning_rate = value
        return _HyperParameter(value, self._owner)

    def __reduce__(self):
        return (float, (float(self),))


class Optimizer:
    _name = "optimizer"

    def __init__(self, learning_rate=0.001, lr=None, name=None, **kwargs):
        if lr is not None:           # legacy spelling used by the reference's tuner tests
            learning_rate = lr
        self._schedule = learning_rate if callable(learning_rate) else None
        self._lr = float(learning_rate(0)) if self._schedule else float(learning_rate)
        self.iterations = 0
        self.name = name or type(self).__name__

    @property
    def learning_rate(self):
        return _HyperParameter(self._lr, self)

    @learning_rate.setter
    def learning_rate(self, v):
        self._lr = float(v)

    lr = learning_rate

    # -- custom training loops (testdata/mnist_example_using_ctl.py:128-129) --------------------
    def _engine_for(self, variables):
        """The data-parallel engine that owns ``variables``: the one ``Model.compile`` + ``fit``
        created for this optimizer, else one built here on first use."""
        eng = getattr(self, "_engine", None)
        ids = tuple(id(v) for v in variables)
        if eng is not None and getattr(self, "_engine_vars", ids) == ids:
            return eng
        import torch
        from ..distribute import get_strategy
        from ..parallel.engine import DPEngine
        from . import models
        st = get_strategy()
        dev = st.device
        for v in variables:
            if v.device != dev:
                v.data = v.data.to(dev)
        cd = torch.bfloat16 if dev.type == "cuda" and models.mixed_precision.global_policy() == "mixed_bfloat16" \
            else torch.float32
        named = [("variable_{}".format(i), v) for i, v in enumerate(variables)]
        self._engine = DPEngine(named, self.spec(), st.communicator, compute_dtype=cd)
        self._engine_vars = ids
        return self._engine

    def apply_gradients(self, grads_and_vars, name=None, **kwargs):
        """Sums the gradients over replicas and applies the update (one fused kernel per bucket).
        As in ``tf.distribute``, per-replica losses are expected to be scaled by 1 / global batch."""
        pairs = [(g, v) for g, v in grads_and_vars]
        variables = [v for _g, v in pairs if v.requires_grad]
        eng = self._engine_for(variables)
        if not eng.in_step:
            eng.begin_step()
        eng.set_lr(self.current_lr())
        import torch
        with torch.no_grad():
            for g, v in pairs:
                if g is not None and v.requires_grad:
                    v.grad.copy_(g)
        eng.end_step()
        self.iterations += 1
        return self.iterations

    def minimize(self, loss, var_list, grad_loss=None, name=None, tape=None):
        """Backward pass with the bucketed all-reduce + update overlapped (the path ``fit`` uses).
        ``loss`` may be a tensor or, as in Keras, a callable returning one."""
        variables = [v for v in (var_list() if callable(var_list) else var_list) if v.requires_grad]
        if getattr(self, "_engine", None) is None or \
                getattr(self, "_engine_vars", None) not in (None, tuple(id(v) for v in variables)):
            # the engine moves the variables into its arena: not possible between forward and backward
            from .tape import GradientTape
            t = tape if tape is not None else GradientTape()
            value = loss() if callable(loss) else loss
            return self.apply_gradients(zip(t.gradient(value, variables), variables))
        eng = self._engine
        if not eng.in_step:
            eng.begin_step()
        eng.set_lr(self.current_lr())
        value = loss() if callable(loss) else loss
        value.backward(grad_loss)
        eng.end_step()
        self.iterations += 1
        return self.iterations

    def current_lr(self) -> float:
        if self._schedule is not None:
            return float(self._schedule(self.iterations))
        return self._lr

    def spec(self) -> OptimizerSpec:
        raise NotImplementedError

    def get_config(self):
        cfg = {"name": self.name, "learning_rate": self._lr}
        if self._schedule is not None:
            cfg["learning_rate"] = {"class_name": type(self._schedule).__name__,
                                    "config": self._schedule.get_config()}
        return cfg


class Adam(Optimizer):
    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, amsgrad=False,
                 **kwargs):
        super().__init__(learning_rate, **kwargs)
        if amsgrad:
            raise NotImplementedError("amsgrad is not supported by the fused Adam kernel")
        self.beta_1, self.beta_2, self.epsilon = beta_1, beta_2, epsilon

    def spec(self):
        return OptimizerSpec("adam", lr=self.current_lr(), beta1=self.beta_1, beta2=self.beta_2,
                             eps=self.epsilon)

    def get_config(self):
        return {**super().get_config(), "beta_1": self.beta_1, "beta_2": self.beta_2,
                "epsilon": self.epsilon}


class AdamW(Adam):
    """Adam with decoupled weight decay (``tf.keras.optimizers.AdamW``, ``torch.optim.AdamW``):
    ``p <- p (1 - lr * weight_decay)`` before the Adam step; the same fused kernel, optimizer code 3."""

    def __init__(self, learning_rate=0.001, weight_decay=0.004, beta_1=0.9, beta_2=0.999, epsilon=1e-7,
                 **kwargs):
        super().__init__(learning_rate, beta_1=beta_1, beta_2=beta_2, epsilon=epsilon, **kwargs)
        self.weight_decay = float(weight_decay)

    def spec(self):
        return OptimizerSpec("adamw", lr=self.current_lr(), beta1=self.beta_1, beta2=self.beta_2,
                             eps=self.epsilon, weight_decay=self.weight_decay)

    def get_config(self):
        return {**super().get_config(), "weight_decay": self.weight_decay}


class SGD(Optimizer):
    def __init__(self, learning_rate=0.01, momentum=0.0, nesterov=False, **kwargs):
        super().__init__(learning_rate, **kwargs)
        self.momentum, self.nesterov = momentum, nesterov

    def spec(self):
        return OptimizerSpec("sgd", lr=self.current_lr(), momentum=self.momentum,
                             nesterov=self.nesterov)

    def get_config(self):
        return {**super().get_config(), "momentum": self.momentum, "nesterov": self.nesterov}


class RMSprop(Optimizer):
    def __init__(self, learning_rate=0.001, rho=0.9, momentum=0.0, epsilon=1e-7, centered=False,
                 **kwargs):
        super().__init__(learning_rate, **kwargs)
        if centered:
            raise NotImplementedError("centered RMSprop is not supported by the fused kernel")
        self.rho, self.momentum, self.epsilon = rho, momentum, epsilon

    def spec(self):
        return OptimizerSpec("rmsprop", lr=self.current_lr(), momentum=self.rho, eps=self.epsilon,
                             rms_momentum=self.momentum)

    def get_config(self):
        return {**super().get_config(), "rho": self.rho, "momentum": self.momentum,
                "epsilon": self.epsilon}


_BY_NAME = {"adam": Adam, "adamw": AdamW, "sgd": SGD, "rmsprop": RMSprop}


def get(identifier):
    if isinstance(identifier, Optimizer):
        return identifier
    if isinstance(identifier, str):
        key = identifier.lower()
        if key in _BY_NAME:
            return _BY_NAME[key]()
    if isinstance(identifier, dict):
        return deserialize(identifier)
    raise ValueError(f"Could not interpret optimizer identifier: {identifier}")


def serialize(opt: Optimizer) -> dict:
    return {"class_name": type(opt).__name__, "config": opt.get_config()}


def deserialize(d: dict) -> Optimizer:
    cls = _BY_NAME[d["class_name"].lower()]
    cfg = dict(d.get("config", {}))
    cfg.pop("name", None)
    lr = cfg.get("learning_rate")
    if isinstance(lr, dict):
        cfg["learning_rate"] = getattr(schedules, lr["class_name"])(**lr["config"])
    return cls(**cfg)
