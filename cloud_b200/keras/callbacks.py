"""Keras callbacks used by the reference's examples and by the tuner.

``LearningRateScheduler`` + ``ModelCheckpoint`` (core/tests/testdata/mnist_example_using_fit.py:95-97,
save_and_load.py:92-96), ``TensorBoard`` + ``EarlyStopping``
(core/tests/examples/call_run_within_script_with_keras_fit.py:93-101).  TensorBoard here writes
scalars as JSON lines (``<log_dir>/{train,validation}/scalars.jsonl``) - the local stand-in for the
hosted TensorBoard the reference points users to (README.md:415-416).
"""
from __future__ import annotations

import json
import os
import time

import numpy as np


class Callback:
    def __init__(self):
        self.model = None
        self.params = {}

    def set_model(self, model):
        self.model = model

    def set_params(self, params):
        self.params = params

    def on_train_begin(self, logs=None): pass
    def on_train_end(self, logs=None): pass
    def on_epoch_begin(self, epoch, logs=None): pass
    def on_epoch_end(self, epoch, logs=None): pass
    def on_train_batch_begin(self, batch, logs=None): pass
    def on_train_batch_end(self, batch, logs=None): pass
    def on_test_begin(self, logs=None): pass
    def on_test_end(self, logs=None): pass
    def on_test_batch_begin(self, batch, logs=None): pass
    def on_test_batch_end(self, batch, logs=None): pass
    def on_predict_begin(self, logs=None): pass
    def on_predict_end(self, logs=None): pass
    # legacy aliases
    def on_batch_begin(self, batch, logs=None): pass
    def on_batch_end(self, batch, logs=None): pass


class CallbackList:
    def __init__(self, callbacks, model, params=None):
        self.callbacks = list(callbacks or [])
        for cb in self.callbacks:
            cb.set_model(model)
            cb.set_params(params or {})

    def needs_batch_logs(self) -> bool:
        for cb in self.callbacks:
            for name in ("on_train_batch_end", "on_batch_end", "on_train_batch_begin", "on_batch_begin"):
                if getattr(type(cb), name) is not getattr(Callback, name):
                    return True
        return False

    def call(self, hook, *args):
        for cb in self.callbacks:
            getattr(cb, hook)(*args)
            if hook == "on_train_batch_end":
                cb.on_batch_end(*args)
            elif hook == "on_train_batch_begin":
                cb.on_batch_begin(*args)


class History(Callback):
    def on_train_begin(self, logs=None):
        self.epoch = []
        self.history = {}

    def on_epoch_end(self, epoch, logs=None):
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(v)


class ProgbarLogger(Callback):
    """verbose=1/2 console output in the Keras format (one line per epoch)."""

    def __init__(self, verbose=1):
        super().__init__()
        self.verbose = verbose

    def on_epoch_begin(self, epoch, logs=None):
        self._t0 = time.time()
        if self.verbose:
            print(f"Epoch {epoch + 1}/{self.params.get('epochs', '?')}", flush=True)

    def on_epoch_end(self, epoch, logs=None):
        if not self.verbose:
            return
        # a dataset of unknown size (generator, filter) has been counted by the end of the epoch
        steps = self.params.get("steps") or (getattr(self.model, "_fit_profile", None) or {}).get("steps") or "?"
        dt = time.time() - self._t0
        body = " - ".join(f"{k}: {v:.4f}" for k, v in (logs or {}).items())
        print(f"{steps}/{steps} - {dt:.0f}s - {body}", flush=True)


class LearningRateScheduler(Callback):
    def __init__(self, schedule, verbose=0):
        super().__init__()
        self.schedule = schedule
        self.verbose = verbose

    def on_epoch_begin(self, epoch, logs=None):
        try:
            lr = self.schedule(epoch, float(self.model.optimizer.learning_rate))
        except TypeError:
            lr = self.schedule(epoch)
        self.model.optimizer.learning_rate = float(lr)
        if self.verbose:
            print(f"\nEpoch {epoch + 1}: LearningRateScheduler setting learning rate to {lr}.")

    def on_epoch_end(self, epoch, logs=None):
        if logs is not None:
            logs["lr"] = float(self.model.optimizer.learning_rate)


def _improved(monitor, mode, current, best):
    if best is None:
        return True
    if mode == "auto":
        mode = "max" if ("acc" in monitor or monitor.startswith("fmeasure") or "auc" in monitor) else "min"
    return current > best if mode == "max" else current < best


class ModelCheckpoint(Callback):
    """Saves the model (or its weights) at epoch end.  Only the chief replica writes; the other
    replicas take part in the gather of sharded state (cloud_fit/remote.py:130-145 semantics)."""

    def __init__(self, filepath, monitor="val_loss", verbose=0, save_best_only=False,
                 save_weights_only=False, mode="auto", save_freq="epoch", **kwargs):
        super().__init__()
        self.filepath = str(filepath)
        self.monitor, self.mode = monitor, mode
        self.verbose = verbose
        self.save_best_only = save_best_only
        self.save_weights_only = save_weights_only
        self.best = None

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        if self.save_best_only:
            cur = logs.get(self.monitor)
            if cur is None or not _improved(self.monitor, self.mode, cur, self.best):
                return
            self.best = cur
        path = self.filepath.format(epoch=epoch + 1, **logs)
        if self.save_weights_only:
            self.model.save_weights(path)
        else:
            self.model.save(path)
        if self.verbose:
            print(f"\nEpoch {epoch + 1}: saving model to {path}")


class EarlyStopping(Callback):
    def __init__(self, monitor="val_loss", min_delta=0, patience=0, verbose=0, mode="auto",
                 baseline=None, restore_best_weights=False):
        super().__init__()
        self.monitor, self.min_delta, self.patience = monitor, abs(min_delta), patience
        self.mode, self.verbose = mode, verbose
        self.restore_best_weights = restore_best_weights
        self.stopped_epoch = 0

    def on_train_begin(self, logs=None):
        self.wait, self.best, self.best_weights = 0, None, None

    def on_epoch_end(self, epoch, logs=None):
        cur = (logs or {}).get(self.monitor)
        if cur is None:
            return
        mode = self.mode
        if mode == "auto":
            mode = "max" if "acc" in self.monitor else "min"
        better = self.best is None or (cur - self.min_delta > self.best if mode == "max"
                                       else cur + self.min_delta < self.best)
        if better:
            self.best, self.wait = cur, 0
            if self.restore_best_weights:
                self.best_weights = self.model.get_weights()
        else:
            self.wait += 1
            if self.wait >= self.patience:
                self.stopped_epoch = epoch
                self.model.stop_training = True
                if self.restore_best_weights and self.best_weights is not None:
                    self.model.set_weights(self.best_weights)

    def on_train_end(self, logs=None):
        if self.stopped_epoch > 0 and self.verbose:
            print(f"Epoch {self.stopped_epoch + 1}: early stopping")


class TensorBoard(Callback):
    """Logs for TensorBoard (``tf.keras.callbacks.TensorBoard``): real event files
    (``<log_dir>/{train,validation}/events.out.tfevents.*``, tags ``epoch_<metric>``, weight
    histograms every ``histogram_freq`` epochs - see ``cloud_b200/io/tfevents.py``), plus the same
    scalars as JSON lines (``scalars.jsonl``) for tools without a protobuf reader.  Chief only."""

    def __init__(self, log_dir="logs", histogram_freq=0, write_graph=True, update_freq="epoch",
                 **kwargs):
        super().__init__()
        self.log_dir = str(log_dir)
        self.histogram_freq = histogram_freq
        self._writers = {}

    def _chief(self):
        return getattr(self.model, "_is_chief", True)

    def _writer(self, split):
        if split not in self._writers:
            from ..io import tfevents
            self._writers[split] = tfevents.EventFileWriter(os.path.join(self.log_dir, split))
        return self._writers[split]

    def _write(self, split, step, scalars):
        if not self._chief():
            return
        d = os.path.join(self.log_dir, split)
        os.makedirs(d, exist_ok=True)
        now = time.time()
        writer = self._writer(split)
        for k, v in scalars.items():
            writer.add_scalar("epoch_" + k, float(v), int(step), now)
        writer.flush()
        with open(os.path.join(d, "scalars.jsonl"), "a") as f:
            f.write(json.dumps({"step": int(step), "wall_time": now,
                                **{k: float(v) for k, v in scalars.items()}}) + "\n")

    def on_train_begin(self, logs=None):
        if self._chief():
            os.makedirs(os.path.join(self.log_dir, "train"), exist_ok=True)
            self._writer("train")

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        train = {k: v for k, v in logs.items() if not k.startswith("val_")}
        val = {k[4:]: v for k, v in logs.items() if k.startswith("val_")}
        if train:
            self._write("train", epoch, train)
        if val:
            self._write("validation", epoch, val)
        if self.histogram_freq and (epoch + 1) % self.histogram_freq == 0 and self._chief():
            hist = {}
            writer = self._writer("train")
            for layer in self.model.layers:
                for name, w in zip(layer.weight_names, layer.get_weights()):
                    counts, edges = np.histogram(w, bins=30)
                    hist[name] = {"counts": counts.tolist(), "edges": edges.tolist()}
                    writer.add_histogram(name.replace(":0", "_0"), w, int(epoch))
            writer.flush()
            with open(os.path.join(self.log_dir, "train", "histograms.jsonl"), "a") as f:
                f.write(json.dumps({"step": int(epoch), "histograms": hist}) + "\n")

    def on_train_end(self, logs=None):
        for w in self._writers.values():
            w.close()
        self._writers = {}

    def __getstate__(self):                 # cloud_fit pickles callbacks: open files stay behind
        state = dict(self.__dict__)
        state["_writers"] = {}
        return state


class LambdaCallback(Callback):
    def __init__(self, on_epoch_begin=None, on_epoch_end=None, on_batch_begin=None,
                 on_batch_end=None, on_train_begin=None, on_train_end=None):
        super().__init__()
        for name, fn in (("on_epoch_begin", on_epoch_begin), ("on_epoch_end", on_epoch_end),
                         ("on_train_batch_begin", on_batch_begin), ("on_train_batch_end", on_batch_end),
                         ("on_train_begin", on_train_begin), ("on_train_end", on_train_end)):
            if fn is not None:
                setattr(self, name, fn)


class TerminateOnNaN(Callback):
    def on_epoch_end(self, epoch, logs=None):
        loss = (logs or {}).get("loss")
        if loss is not None and (np.isnan(loss) or np.isinf(loss)):
            self.model.stop_training = True


class ReduceLROnPlateau(Callback):
    """Multiplies the learning rate by ``factor`` when ``monitor`` has not improved for
    ``patience`` epochs (``tf.keras.callbacks.ReduceLROnPlateau``)."""

    def __init__(self, monitor="val_loss", factor=0.1, patience=10, verbose=0, mode="auto",
                 min_delta=1e-4, cooldown=0, min_lr=0.0, **kwargs):
        super().__init__()
        if factor >= 1.0:
            raise ValueError("ReduceLROnPlateau does not support a factor >= 1.0.")
        self.monitor, self.factor, self.patience, self.verbose = monitor, factor, patience, verbose
        self.mode, self.min_delta, self.cooldown, self.min_lr = mode, abs(min_delta), cooldown, min_lr

    def on_train_begin(self, logs=None):
        self.best, self.wait, self.cooldown_counter = None, 0, 0

    def on_epoch_end(self, epoch, logs=None):
        logs = logs if logs is not None else {}
        logs["lr"] = float(self.model.optimizer.learning_rate)
        cur = logs.get(self.monitor)
        if cur is None:
            return
        mode = self.mode if self.mode != "auto" else ("max" if "acc" in self.monitor else "min")
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.wait = 0
        better = self.best is None or (cur > self.best + self.min_delta if mode == "max"
                                       else cur < self.best - self.min_delta)
        if better:
            self.best, self.wait = cur, 0
        elif self.cooldown_counter <= 0:
            self.wait += 1
            if self.wait >= self.patience:
                old = float(self.model.optimizer.learning_rate)
                if old > self.min_lr:
                    new = max(old * self.factor, self.min_lr)
                    self.model.optimizer.learning_rate = new
                    if self.verbose:
                        print(f"\nEpoch {epoch + 1}: ReduceLROnPlateau reducing learning rate to {new}.")
                    self.cooldown_counter, self.wait = self.cooldown, 0


class CSVLogger(Callback):
    """One row of metrics per epoch (``tf.keras.callbacks.CSVLogger``); written by the chief only."""

    def __init__(self, filename, separator=",", append=False):
        super().__init__()
        self.filename, self.sep, self.append = str(filename), separator, append
        self._file, self._keys = None, None

    def _chief(self):
        return getattr(self.model, "_is_chief", True)

    def on_train_begin(self, logs=None):
        if not self._chief():
            return
        import os
        os.makedirs(os.path.dirname(os.path.abspath(self.filename)), exist_ok=True)
        had_rows = self.append and os.path.exists(self.filename) and os.path.getsize(self.filename) > 0
        self._file = open(self.filename, "a" if self.append else "w")
        self._write_header = not had_rows

    def on_epoch_end(self, epoch, logs=None):
        if self._file is None:
            return
        logs = logs or {}
        if self._keys is None:
            self._keys = sorted(logs)
            if self._write_header:
                self._file.write(self.sep.join(["epoch"] + self._keys) + "\n")
        self._file.write(self.sep.join([str(epoch)] + [repr(float(logs[k])) if k in logs else "NA"
                                                        for k in self._keys]) + "\n")
        self._file.flush()

    def on_train_end(self, logs=None):
        if self._file is not None:
            self._file.close()
            self._file = None


class BackupAndRestore(Callback):
    """Fault tolerance for ``fit`` (``tf.keras.callbacks.experimental.BackupAndRestore``): after
    every epoch the training state (weights, optimizer state, epoch number) is saved under
    ``backup_dir``; a ``fit`` that finds a backup there - the same script started again after a
    crash, e.g. by the launcher with ``TFC_JOB_RESTARTS`` - restores it and continues with the next
    epoch.  The backup is deleted when training ends normally.  All replicas take part in saving
    (optimizer state may be sharded); only the chief writes."""

    def __init__(self, backup_dir, delete_checkpoint=True):
        super().__init__()
        self.backup_dir = str(backup_dir)
        self.delete_checkpoint = delete_checkpoint

    def _meta(self):
        import os
        return os.path.join(self.backup_dir, "backup.json")

    def on_train_begin(self, logs=None):
        import json
        import os
        try:
            with open(self._meta()) as f:
                meta = json.load(f)
        except (OSError, ValueError):
            return
        from . import saving
        ckpt = os.path.join(self.backup_dir, meta["checkpoint"])
        saving.restore_training_state(self.model, ckpt)
        self.model._resume_from_epoch = int(meta["epoch"]) + 1
        if getattr(self.model, "_is_chief", True):
            print("BackupAndRestore: resuming after epoch {} from {}".format(int(meta["epoch"]) + 1, ckpt), flush=True)

    def on_epoch_end(self, epoch, logs=None):
        import json
        import os
        import shutil
        name = "ckpt_{}".format(epoch % 2)                 # two slots: the previous one stays valid
        path = os.path.join(self.backup_dir, name)
        chief = getattr(self.model, "_is_chief", True)
        if chief:
            shutil.rmtree(path, ignore_errors=True)
        self.model.save(path)                              # collective
        if chief:
            tmp = self._meta() + ".tmp"
            with open(tmp, "w") as f:
                json.dump({"epoch": epoch, "checkpoint": name}, f)
            os.replace(tmp, self._meta())                  # the backup exists from this instant

    def on_train_end(self, logs=None):
        import shutil
        if self.delete_checkpoint and getattr(self.model, "_is_chief", True) \
                and not getattr(self.model, "stop_training_failed", False):
            shutil.rmtree(self.backup_dir, ignore_errors=True)


class experimental:  # tf.keras.callbacks.experimental.*
    BackupAndRestore = BackupAndRestore
