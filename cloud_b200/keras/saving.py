"""Saving / loading in Keras-compatible layouts.

* ``save_weights('x.h5')`` writes the Keras HDF5 weight layout (root attrs ``layer_names``,
  ``backend``, ``keras_version``; one group per layer with ``weight_names``; datasets
  ``<layer>/<layer>/kernel:0`` with Dense kernels ``[in,out]`` and Conv kernels ``HWIO``) so that
  the reference side (Keras ``load_weights``) can read it - SURVEY.md appendix B4.
* ``save_weights('ckpt/cp.ckpt')`` (TF-checkpoint spelling, core/tests/testdata/save_and_load.py:92-96)
  writes the same HDF5 content to ``<prefix>.weights.h5`` plus a ``checkpoint`` marker file.
* ``model.save(dir)`` writes ``config.json`` + ``weights.h5`` + ``optimizer.pt`` (SavedModel-like
  directory, what cloud_fit ships around: experimental/cloud_fit/client.py:192, remote.py:145);
  ``model.save('m.h5')`` writes one HDF5 file with ``model_config`` / ``training_config`` attrs and
  a ``model_weights`` group.
* Only the chief replica writes; every replica takes part in gathering sharded optimizer state.
  BatchNorm moving statistics are averaged across replicas first (MEAN aggregation on read).
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch

from ..io import hdf5
from ..version import __version__

KERAS_VERSION = b"2.4.0"
BACKEND = b"tensorflow"


def _is_h5(path: str) -> bool:
    return str(path).endswith((".h5", ".hdf5", ".keras"))


def _weights_path(path: str) -> str:
    path = str(path)
    return path if _is_h5(path) else path + ".weights.h5"


def _sync_replica_stats(model) -> None:
    eng = getattr(model, "_engine", None)
    if eng is None or eng.world == 1:
        return
    for buf in model.buffers():
        if buf.is_floating_point():
            t = buf.detach().clone().float()
            eng.comm.all_reduce_sum(t)
            buf.copy_((t / eng.world).to(buf.dtype))


def _write_weight_groups(group, model) -> None:
    layers = [l for l in model._layers if not isinstance(l, _input_layer_cls())]
    flat = []
    for l in layers:
        flat.append(l)
    group.attrs["layer_names"] = [l.name.encode("utf8") for l in flat]
    group.attrs["backend"] = BACKEND
    group.attrs["keras_version"] = KERAS_VERSION
    for l in flat:
        g = group.create_group(l.name)
        if hasattr(l, "_weight_layers"):        # nested model: Keras flattens its weights here
            names, values = [], []
            for sub in l._weight_layers():
                names.extend(sub.weight_names)
                values.extend(sub.get_weights())
        else:
            names = list(l.weight_names) if getattr(l, "_keras_weight_specs", None) else []
            values = l.get_weights() if names else []
        g.attrs["weight_names"] = [n.encode("utf8") for n in names] if names else np.zeros((0,), "S1")
        for n, v in zip(names, values):
            g.create_dataset(n, data=np.asarray(v, dtype=np.float32))


def _input_layer_cls():
    from .layers import InputLayer
    return InputLayer


def save_weights(model, filepath) -> None:
    _sync_replica_stats(model)
    if not model._is_chief:
        return
    path = _weights_path(filepath)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with hdf5.File(path, "w") as f:
        _write_weight_groups(f, model)
    if not _is_h5(str(filepath)):
        with open(os.path.join(os.path.dirname(os.path.abspath(path)), "checkpoint"), "w") as fh:
            fh.write(f'model_checkpoint_path: "{os.path.basename(str(filepath))}"\n')


def _read_weight_groups(group, model, by_name=False) -> None:
    names = [n.decode("utf8") if isinstance(n, bytes) else str(n) for n in group.attrs["layer_names"]]
    layers = [l for l in model._layers if not isinstance(l, _input_layer_cls())]
    by = {l.name: l for l in layers}
    if not by_name:
        file_layers = [n for n in names if len(group[n].attrs["weight_names"])]
        model_layers = [l for l in layers if _n_weights(l)]
        if len(file_layers) != len(model_layers):
            raise ValueError(f"You are trying to load a weight file containing {len(file_layers)} "
                             f"layers into a model with {len(model_layers)} layers.")
        pairs = list(zip(file_layers, model_layers))
    else:
        pairs = [(n, by[n]) for n in names if n in by and _n_weights(by[n])]
    for fname, layer in pairs:
        g = group[fname]
        wnames = [n.decode("utf8") if isinstance(n, bytes) else str(n) for n in g.attrs["weight_names"]]
        values = [g[n] for n in wnames]
        if hasattr(layer, "_weight_layers"):
            i = 0
            for sub in layer._weight_layers():
                k = len(sub._keras_weight_specs)
                sub.set_weights(values[i:i + k])
                i += k
        else:
            layer.set_weights(values)
    if getattr(model, "_engine", None) is not None:
        model._engine.params_changed()


def _n_weights(layer) -> int:
    if hasattr(layer, "_weight_layers"):
        return sum(len(s._keras_weight_specs) for s in layer._weight_layers())
    return len(getattr(layer, "_keras_weight_specs", []) or [])


def load_weights(model, filepath, by_name=False) -> None:
    path = str(filepath)
    if not _is_h5(path) and os.path.exists(_weights_path(path)):
        path = _weights_path(path)        # TF-checkpoint style prefix -> "<prefix>.weights.h5"
    elif not os.path.exists(path) or os.path.isdir(path):
        alt = _weights_path(path)
        if os.path.exists(alt):
            path = alt
        elif os.path.isdir(path) and os.path.exists(os.path.join(path, "weights.h5")):
            path = os.path.join(path, "weights.h5")
        else:
            raise OSError(f"Unable to open file (unable to open file: name = '{filepath}')")
    if not model.built or not all(getattr(l, "built", True) for l in model._layers):
        raise ValueError("Unable to load weights saved in HDF5 format into a subclassed Model "
                         "which has not created its variables yet. Call the Model first, then "
                         "load the weights.")
    with hdf5.File(path, "r") as f:
        root = f["model_weights"] if "model_weights" in f.keys() else f
        _read_weight_groups(root, model, by_name=by_name)


def latest_checkpoint(checkpoint_dir):
    """``tf.train.latest_checkpoint`` (core/tests/testdata/save_and_load.py:103)."""
    marker = os.path.join(str(checkpoint_dir), "checkpoint")
    if not os.path.exists(marker):
        return None
    with open(marker) as fh:
        line = fh.readline()
    name = line.split(":", 1)[1].strip().strip('"')
    return os.path.join(str(checkpoint_dir), name)


# ------------------------------------------------------------------------------------------------
def _training_config(model) -> dict | None:
    if model.optimizer is None:
        return None
    from . import losses as losses_mod, optimizers as opt_mod
    return {"optimizer_config": opt_mod.serialize(model.optimizer),
            "loss": losses_mod.serialize(model.loss) if model.loss is not None else None,
            "metrics": [m.name for m in model._metrics]}


def save_model(model, filepath, include_optimizer=True) -> None:
    _sync_replica_stats(model)
    eng = getattr(model, "_engine", None)
    opt_state = None
    if include_optimizer and eng is not None:
        opt_state = eng.full_master_state()      # collective: every replica participates
    if not model._is_chief:
        return
    path = str(filepath)
    config = {"class_name": type(model).__name__, "config": model.get_config(),
              "cloud_b200_version": __version__}
    train_cfg = _training_config(model)
    if _is_h5(path):
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        with hdf5.File(path, "w") as f:
            blob = json.dumps(config).encode("utf8")
            try:
                f.attrs["model_config"] = blob
            except hdf5.AttributeTooLarge:
                f.create_dataset("model_config_json", data=np.frombuffer(blob, dtype=np.uint8))
            if train_cfg is not None:
                f.attrs["training_config"] = json.dumps(train_cfg).encode("utf8")
            f.attrs["backend"] = BACKEND
            f.attrs["keras_version"] = KERAS_VERSION
            _write_weight_groups(f.create_group("model_weights"), model)
        if opt_state is not None:
            torch.save(opt_state, path + ".optimizer.pt")
        return
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as fh:
        json.dump({"model": config, "training": train_cfg}, fh)
    with hdf5.File(os.path.join(path, "weights.h5"), "w") as f:
        _write_weight_groups(f, model)
    if opt_state is not None:
        torch.save(opt_state, os.path.join(path, "optimizer.pt"))
    with open(os.path.join(path, "saved_model.json"), "w") as fh:   # marker, like saved_model.pb
        json.dump({"format": "cloud_b200", "version": __version__}, fh)


def load_model(filepath, custom_objects=None, compile=True):
    from . import losses as losses_mod, optimizers as opt_mod
    from .models import Model, Sequential
    path = str(filepath)
    opt_path = None
    if _is_h5(path):
        with hdf5.File(path, "r") as f:
            if "model_config" in f.attrs:
                blob = f.attrs["model_config"]
            elif "model_config_json" in f.keys():
                blob = f["model_config_json"].tobytes()
            else:
                raise ValueError("No model config found in the file at " + path)
            config = json.loads(blob.decode("utf8") if isinstance(blob, bytes) else blob)
            train_cfg = json.loads(f.attrs["training_config"].decode("utf8")) \
                if "training_config" in f.attrs else None
            model = _model_from_config(config, custom_objects, Model, Sequential)
            _read_weight_groups(f["model_weights"], model)
        opt_path = path + ".optimizer.pt"
    else:
        cfg_file = os.path.join(path, "config.json")
        if not os.path.exists(cfg_file):
            raise OSError(f"SavedModel file does not exist at: {path}/{{saved_model.pbtxt|saved_model.pb}}")
        with open(cfg_file) as fh:
            blob = json.load(fh)
        config, train_cfg = blob["model"], blob.get("training")
        model = _model_from_config(config, custom_objects, Model, Sequential)
        with hdf5.File(os.path.join(path, "weights.h5"), "r") as f:
            _read_weight_groups(f, model)
        opt_path = os.path.join(path, "optimizer.pt")
    if compile and train_cfg:
        model.compile(optimizer=opt_mod.deserialize(train_cfg["optimizer_config"]),
                      loss=losses_mod.deserialize(train_cfg["loss"]) if train_cfg.get("loss") else None,
                      metrics=train_cfg.get("metrics") or [])
        if opt_path and os.path.exists(opt_path):
            model._pending_optimizer_state = torch.load(opt_path, weights_only=False)
    return model


def restore_training_state(model, filepath) -> None:
    """Weights and optimizer state of a ``model.save(dir)`` into an existing, compiled model of the
    same architecture (what ``BackupAndRestore`` does after a restart)."""
    path = str(filepath)
    model.load_weights(os.path.join(path, "weights.h5"))
    opt_path = os.path.join(path, "optimizer.pt")
    if not os.path.exists(opt_path):
        return
    state = torch.load(opt_path, weights_only=False)
    eng = getattr(model, "_engine", None)
    if eng is not None:
        eng.load_master_state(state)
    else:
        model._pending_optimizer_state = state
    if model.optimizer is not None and "__step__" in state:
        model.optimizer.iterations = int(state["__step__"])


def _model_from_config(config, custom_objects, Model, Sequential):
    cfg = config["config"]
    cls = Sequential if (config.get("class_name") == "Sequential" or cfg.get("kind") == "Sequential") \
        else Model
    return cls.from_config(cfg, custom_objects)
