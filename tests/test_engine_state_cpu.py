"""Engine state that checkpoints depend on: the fp32 master - not its bf16
rounding - is what ``get_weights`` / ``save_weights`` see and what ``set_weights`` / ``load_weights``
restore; optimizer state is stored in the parameter's logical shape; the tail batch of an epoch is
split so that no replica is left without rows."""
import numpy as np
import pytest
import torch

import cloud_b200 as tfc
from cloud_b200.keras import layers
from cloud_b200.keras.batch_sources import _ArraySource
from cloud_b200.parallel.comm import Communicator
from cloud_b200.parallel.engine import DPEngine, OptimizerSpec

keras = tfc.keras


@pytest.mark.parametrize("n,batch,world", [(64 * 3 + r, 64, 8) for r in range(0, 9)] + [(70, 32, 4), (10, 4, 3)])
def test_tail_batch_split_leaves_no_replica_empty(n, batch, world):
    x = np.arange(n, dtype=np.float32).reshape(n, 1)
    y = np.arange(n)
    per_rank = [list(_ArraySource(x, y, batch, False, world, r)) for r in range(world)]
    steps = {len(p) for p in per_rank}
    assert len(steps) == 1                                  # every replica runs the same number of steps
    for s in range(steps.pop()):
        rows = [p[s][0] for p in per_rank]
        assert all(len(r) > 0 for r in rows), [len(r) for r in rows]
        assert max(len(r) for r in rows) - min(len(r) for r in rows) <= 1
        got = torch.cat(rows).reshape(-1)
        lo = s * batch
        assert torch.equal(got, torch.arange(lo, lo + len(got), dtype=torch.float32))   # nothing lost or repeated
        assert int(per_rank[0][s][2]) == len(got)           # the global batch size every replica divides by
    # rows that are dropped: only a tail shorter than the number of replicas
    total = sum(len(item[0]) for p in per_rank for item in p)
    assert n - total < world


def _bf16_engine(params):
    comm = Communicator(torch.device("cpu"), single=True)
    return DPEngine(params, OptimizerSpec(name="adam", lr=1e-2), comm, compute_dtype=torch.bfloat16)


def test_fp32_master_is_what_get_and_set_weights_see():
    torch.manual_seed(0)
    lin = torch.nn.Linear(16, 8)
    conv = torch.nn.Conv2d(4, 6, 3)
    named = [("lin.w", lin.weight), ("lin.b", lin.bias), ("conv.w", conv.weight)]
    eng = _bf16_engine(named)
    assert lin.weight.dtype == torch.bfloat16
    # a value bf16 cannot hold: 1 + 2^-12
    exact = torch.full((8, 16), 1.0 + 2.0 ** -12)
    lin.weight._cb2_fp32.set(exact)
    eng.params_changed()                                    # what load_weights calls afterwards
    assert torch.equal(lin.weight._cb2_fp32.get(), exact)   # master kept the 2^-12
    assert torch.equal(lin.weight.detach().float(), exact.to(torch.bfloat16).float())   # arena holds the rounding
    # a conv kernel comes back in its logical O,I,H,W shape
    cw = torch.randn(6, 4, 3, 3)
    conv.weight._cb2_fp32.set(cw)
    assert torch.equal(conv.weight._cb2_fp32.get(), cw)
    # one optimizer step moves the master; the accessor follows it
    eng.begin_step()
    lin.weight.grad.fill_(1.0)
    for _n, p in named:
        p._cb2_grad_written(p)
    eng.end_step()
    after = lin.weight._cb2_fp32.get()
    assert (after < exact).all() and after.dtype == torch.float32
    assert not torch.equal(after, after.to(torch.bfloat16).float())        # still finer than bf16


def test_optimizer_state_is_stored_in_logical_shape_and_restored():
    torch.manual_seed(1)
    conv = torch.nn.Conv2d(4, 6, 3)
    eng = _bf16_engine([("conv.w", conv.weight), ("conv.b", conv.bias)])
    eng.begin_step()
    conv.weight.grad.copy_(torch.randn(6, 4, 3, 3).to(torch.bfloat16))
    conv.bias.grad.fill_(0.5)
    for p in (conv.weight, conv.bias):
        p._cb2_grad_written(p)
    eng.end_step()
    state = eng.full_master_state()
    assert state["__layout__"] == "logical"
    assert tuple(state["conv.w"]["master"].shape) == (6, 4, 3, 3)
    assert torch.allclose(state["conv.w"]["master"], conv.weight.detach().float(), atol=1e-2)
    conv2 = torch.nn.Conv2d(4, 6, 3)
    eng2 = _bf16_engine([("conv.w", conv2.weight), ("conv.b", conv2.bias)])
    eng2.load_master_state(state)
    again = eng2.full_master_state()
    for key in ("master", "s1", "s2"):
        assert torch.equal(again["conv.w"][key], state["conv.w"][key])
    assert again["__step__"] == state["__step__"]


def test_recompile_drops_every_captured_graph_and_the_old_engine_hooks():
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    keras.backend.clear_session()
    m = keras.Sequential([layers.Dense(8, activation="relu", input_shape=(4,)), layers.Dense(2, activation="softmax")])
    m.compile(optimizer="adam", loss="sparse_categorical_crossentropy", steps_per_execution=2)
    x, y = np.random.rand(16, 4).astype("float32"), np.random.randint(0, 2, 16)
    m.fit(x, y, batch_size=4, epochs=1, verbose=0)
    old = m._engine
    m._graph_k = {"key": "stale", "graph": object()}        # what a CUDA run would have left behind
    m.compile(optimizer="adam", loss="sparse_categorical_crossentropy", steps_per_execution=2)
    assert m._engine is None and m._graph is None and m._graph_k is None
    assert old._hook_handles == []
    w0 = [w.copy() for w in m.get_weights()]
    m.fit(x, y, batch_size=4, epochs=1, verbose=0)
    assert m._engine is not old
    assert any(not np.array_equal(a, b) for a, b in zip(w0, m.get_weights()))   # the live weights train


def test_conv3x3_routing_rules(monkeypatch):
    """Which 3x3 layers go to the implicit-GEMM kernel is a pure function of shapes, layout and the
    TFC_CONV3X3* switches (ops/conv.py::conv3x3_own) - checked here without a GPU."""
    import torch
    from cloud_b200.ops import conv as conv_mod

    def xw(width, cin, cout, k=3, channels_last=True):
        x = torch.zeros(1, cin, width, width, dtype=torch.bfloat16)
        w = torch.zeros(cout, cin, k, k, dtype=torch.bfloat16)
        if channels_last:
            w = w.contiguous(memory_format=torch.channels_last)
        return x, w

    monkeypatch.setattr(conv_mod, "_CONV3", "1")
    monkeypatch.setattr(conv_mod, "_CONV3_MIN_W", 14)
    monkeypatch.setattr(conv_mod, "_CONV3_MIN_C", 64)
    assert conv_mod.conv3x3_own(*xw(56, 64, 64))
    assert conv_mod.conv3x3_own(*xw(28, 128, 128))
    assert conv_mod.conv3x3_own(*xw(14, 256, 256))
    assert not conv_mod.conv3x3_own(*xw(7, 512, 512))            # 7x7 stage: mostly padding, stays on cuDNN
    assert not conv_mod.conv3x3_own(*xw(64, 64, 64))             # wider than one strip row allows
    assert not conv_mod.conv3x3_own(*xw(28, 128, 128, k=5))      # only 3x3
    assert not conv_mod.conv3x3_own(*xw(28, 128, 128, channels_last=False))   # weights must be [O, H, W, I]
    monkeypatch.setattr(conv_mod, "_CONV3_MIN_C", 128)
    assert not conv_mod.conv3x3_own(*xw(56, 64, 64))
    monkeypatch.setattr(conv_mod, "_CONV3", "all")
    assert conv_mod.conv3x3_own(*xw(7, 512, 512))
    monkeypatch.setattr(conv_mod, "_CONV3", "0")
    assert not conv_mod.conv3x3_own(*xw(28, 128, 128))
