"""TensorBoard event files: framing and checksums against known CRC-32C vectors, a write / read
round trip, corruption detection, and what the ``TensorBoard`` callback leaves in its log directory
(reference usage: core/tests/examples/call_run_within_script_with_keras_fit.py:94-99)."""
import os
import struct

import numpy as np
import pytest
import torch

import cloud_b200 as tfc
from cloud_b200.io import tfevents

keras = tfc.keras


def test_crc32c_known_vectors():
    assert tfevents.crc32c(b"123456789") == 0xE3069283            # the standard check value
    assert tfevents.crc32c(b"") == 0
    assert tfevents.crc32c(bytes(32)) == 0x8A9136AA               # RFC 3720 B.4: 32 zero bytes
    assert tfevents.crc32c(bytes([0xFF] * 32)) == 0x62A8AB43      # RFC 3720 B.4: 32 0xFF bytes
    c = tfevents.crc32c(b"abc")
    assert tfevents.masked_crc32c(b"abc") == ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


def test_event_file_round_trip(tmp_path):
    w = tfevents.EventFileWriter(str(tmp_path))
    w.add_scalar("epoch_loss", 0.25, 3, wall_time=1234.5)
    w.add_scalar("epoch_accuracy", 0.875, 3)
    values = np.array([-1.0, 0.0, 0.5, 2.0, 2.0])
    w.add_histogram("dense/kernel_0", values, 7, bins=4)
    w.close()
    (path,) = tfevents.event_files(str(tmp_path))
    assert os.path.basename(path).startswith("events.out.tfevents.")
    events = tfevents.read_events(path)
    assert events[0]["file_version"] == "brain.Event:2" and events[0]["step"] == 0
    assert events[1]["scalars"] == {"epoch_loss": 0.25} and events[1]["step"] == 3 and events[1]["wall_time"] == 1234.5
    assert events[2]["scalars"] == {"epoch_accuracy": 0.875}
    h = events[3]["histograms"]["dense/kernel_0"]
    assert (h["min"], h["max"], h["num"], h["sum"], h["sum_squares"]) == (-1.0, 2.0, 5.0, 3.5, 9.25)
    assert h["bucket"] == [1.0, 1.0, 1.0, 2.0] and h["bucket_limit"] == [-0.25, 0.5, 1.25, 2.0]
    # the framing itself: 8-byte little-endian length in front of every record
    raw = open(path, "rb").read()
    (n,) = struct.unpack("<Q", raw[:8])
    assert raw[12:12 + n].endswith(b"brain.Event:2")


def test_corruption_is_detected(tmp_path):
    w = tfevents.EventFileWriter(str(tmp_path))
    w.add_scalar("x", 1.0, 1)
    w.close()
    (path,) = tfevents.event_files(str(tmp_path))
    raw = bytearray(open(path, "rb").read())
    raw[-6] ^= 0x01                                   # a bit inside the last record's payload
    open(path, "wb").write(bytes(raw))
    with pytest.raises(ValueError, match="corrupted"):
        tfevents.read_events(path)
    open(path, "wb").write(bytes(raw[:5]))
    with pytest.raises(ValueError, match="truncated"):
        tfevents.read_events(path)


def test_tensorboard_callback_writes_event_files(tmp_path):
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("/cpu:0"))
    torch.manual_seed(0)
    model = keras.Sequential([keras.layers.Dense(8, activation="relu", input_shape=(4,), name="hidden"),
                              keras.layers.Dense(2, activation="softmax", name="out")])
    model.compile("adam", "sparse_categorical_crossentropy", metrics=["accuracy"])
    x = np.random.rand(64, 4).astype("float32")
    y = (x[:, 0] > 0.5).astype("int64")
    log_dir = tmp_path / "logs"
    hist = model.fit(x, y, batch_size=16, epochs=3, verbose=0, validation_data=(x, y),
                     callbacks=[keras.callbacks.TensorBoard(log_dir=str(log_dir), histogram_freq=1)])
    train = tfevents.read_events(tfevents.event_files(str(log_dir / "train"))[0])
    val = tfevents.read_events(tfevents.event_files(str(log_dir / "validation"))[0])
    losses = [e["scalars"]["epoch_loss"] for e in train if "epoch_loss" in e["scalars"]]
    np.testing.assert_allclose(losses, hist.history["loss"], rtol=1e-6)
    assert [e["step"] for e in train if "epoch_accuracy" in e["scalars"]] == [0, 1, 2]
    val_losses = [e["scalars"]["epoch_loss"] for e in val if "epoch_loss" in e["scalars"]]
    np.testing.assert_allclose(val_losses, hist.history["val_loss"], rtol=1e-6)
    tags = {t for e in train for t in e["histograms"]}
    assert tags == {"hidden/kernel_0", "hidden/bias_0", "out/kernel_0", "out/bias_0"}
    kernel = [e["histograms"]["hidden/kernel_0"] for e in train if "hidden/kernel_0" in e["histograms"]][-1]
    assert kernel["num"] == 32.0 and sum(kernel["bucket"]) == 32.0
    assert (log_dir / "train" / "scalars.jsonl").exists()          # the JSON lines are still there
    import pickle
    pickle.loads(pickle.dumps(keras.callbacks.TensorBoard(log_dir=str(log_dir))))    # cloud_fit pickles callbacks
