"""Port of experimental/cloud_fit/tests/unit/{client,remote,utils}_test.py: the serialisation
round trip (model + datasets + cloudpickled callbacks + fit kwargs), job-spec handling, strategy
validation, and `remote.run` training for real on CPU with a faked TF_CONFIG (remote_test.py:75-121)."""
import json
import os
import pickle
from unittest import mock

import numpy as np
import pytest

import cloud_b200 as tfc
from cloud_b200.core import machine_config
from cloud_b200.experimental.cloud_fit import client, remote, utils
from cloud_b200.launch import registry

keras = tfc.keras
CFG = machine_config.COMMON_MACHINE_CONFIGS


@pytest.fixture(autouse=True)
def _env(monkeypatch, tmp_path):
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    monkeypatch.delenv("TF_CONFIG", raising=False)
    keras.backend.clear_session()


def _model():
    # y = w x + 1 (client_test.py:69-90)
    m = keras.Sequential([keras.layers.Dense(1, input_shape=(1,),
                                             kernel_initializer=keras.initializers.Constant(0.5))])
    m.compile(optimizer=keras.optimizers.SGD(0.002), loss="mse", metrics=["mae"])
    return m


def _xy(n=64):
    x = np.linspace(-1, 1, n, dtype="float32").reshape(-1, 1)
    return x, 3 * x + 1


class CountingCallback(keras.callbacks.Callback):
    def __init__(self, path):
        super().__init__()
        self.path = path

    def on_train_begin(self, logs=None):
        with open(self.path, "a") as f:
            f.write("called\n")


def test_default_job_spec_and_job_id(monkeypatch):
    monkeypatch.setenv("TFC_FAKE_GPUS", "8")
    spec = client._default_job_spec()
    assert spec["worker_count"] == 1 and spec["chief_config"].accelerator_count == 1
    assert client._default_job_id().startswith("cloud_fit_")
    monkeypatch.setenv("TFC_FAKE_GPUS", "0")
    assert client._default_job_spec()["chief_config"] is CFG["CPU"]


def test_serialize_assets_round_trip(tmp_path):
    x, y = _xy()
    ds = tfc.data.Dataset.from_tensor_slices((x, y)).batch(16)
    marker = str(tmp_path / "marker")
    client._serialize_assets(str(tmp_path / "rd"), _model(), x=ds, validation_data=(x, y), epochs=2,
                             batch_size=8, callbacks=[CountingCallback(marker),
                                                      keras.callbacks.TensorBoard(str(tmp_path / "logs"))])
    assets = tmp_path / "rd" / "training_assets"
    assert sorted(os.listdir(tmp_path / "rd")) == ["model", "training_assets"]
    assert (assets / "MANIFEST").read_text().split() == ["batch_size", "callbacks", "epochs",
                                                         "validation_data", "x"]
    kwargs = pickle.loads((assets / "fit_kwargs.pkl").read_bytes())
    assert kwargs["epochs"] == 2 and kwargs["batch_size"] == 8
    xb, yb = next(iter(kwargs["x"]))
    assert xb.shape == (16, 1) and np.allclose(yb, 3 * xb + 1)
    cbs = pickle.loads((assets / "callbacks.pkl").read_bytes())
    assert type(cbs[0]).__name__ == "CountingCallback" and cbs[0].path == marker
    restored = keras.models.load_model(str(tmp_path / "rd" / "model"))
    assert np.allclose(restored.get_weights()[0], 0.5)


def test_generators_are_rejected(tmp_path):
    def gen():
        yield 1
    with pytest.raises(NotImplementedError, match=r"Generators are not supported"):
        client._serialize_assets(str(tmp_path / "rd"), _model(), x=gen())


def test_unsupported_strategy(tmp_path):
    with pytest.raises(ValueError, match=r"is not supported. Must be one of"):
        client.cloud_fit(_model(), str(tmp_path / "rd"), distribution_strategy="OneDeviceStrategy")
    with pytest.raises(ValueError, match=r"is not supported. Must be one of"):
        remote.run(str(tmp_path / "rd"), "TPUStrategy")
    assert set(utils.SUPPORTED_DISTRIBUTION_STRATEGIES) == {"MultiWorkerMirroredStrategy", "MirroredStrategy"}
    assert utils.is_tf_v1() is False


def test_submit_uses_job_spec_job_id_and_args(tmp_path, monkeypatch):
    x, y = _xy()
    calls = {}

    def fake_submit(plan, image, args=None, job_id=None, labels=None, region=None, env=None):
        calls.update(plan=plan, image=image, args=args, job_id=job_id, labels=labels, region=region)
        return job_id
    from cloud_b200.core import deploy
    monkeypatch.setattr(deploy, "submit_plan", fake_submit)
    monkeypatch.setenv("TFC_FAKE_GPUS", "4")
    spec = {"chief_config": CFG["K80_1X"], "worker_count": 2, "worker_config": CFG["K80_1X"],
            "labels": {"k": "v"}}
    rd = str(tmp_path / "rd")
    got = client.cloud_fit(_model(), rd, region="us-east1", job_spec=spec, job_id="my_job",
                           distribution_strategy="MirroredStrategy", x=x, y=y, epochs=1)
    assert got == "my_job" and calls["job_id"] == "my_job" and calls["labels"] == {"k": "v"}
    assert calls["args"] == ["--remote_dir", rd, "--distribution_strategy", "MirroredStrategy"]
    assert calls["plan"]["world_size"] == 3 and calls["region"] == "us-east1"
    assert calls["image"]["entrypoint"][-1] == "cloud_b200.experimental.cloud_fit.remote"
    # CAIP-style spec: replicas are counted
    client.cloud_fit(_model(), rd, job_spec={"trainingInput": {"workerCount": "1"}}, x=x, y=y)
    assert calls["plan"]["world_size"] == 2


def test_remote_run_trains_locally_with_faked_tf_config(tmp_path, monkeypatch):
    x, y = _xy()
    rd = str(tmp_path / "rd")
    marker = str(tmp_path / "marker")
    client._serialize_assets(rd, _model(), x=x, y=y, epochs=30, batch_size=16, verbose=0,
                             callbacks=[CountingCallback(marker),
                                        keras.callbacks.TensorBoard(os.path.join(rd, "logs"))])
    monkeypatch.setenv("TF_CONFIG", json.dumps({
        "cluster": {"worker": ["localhost:9999", "localhost:9999"]},
        "task": {"type": "worker", "index": 0}}))
    trained = remote.run(rd, "MirroredStrategy")
    assert remote._is_current_worker_chief()
    out = keras.models.load_model(os.path.join(rd, "output"))
    before = _model().evaluate(x, y, verbose=0)[0]
    after = out.evaluate(x, y, verbose=0)[0]
    assert after < before * 0.8
    assert np.allclose(out.get_weights()[0], trained.get_weights()[0])
    assert os.path.exists(os.path.join(rd, "logs", "train", "scalars.jsonl"))
    assert open(marker).read() == "called\n"          # the custom callback fired exactly once


def test_remote_main_requires_both_flags():
    with pytest.raises(SystemExit):
        remote.main(["--remote_dir", "/tmp/x"])


def test_cloud_fit_end_to_end_job(tmp_path):
    x, y = _xy()
    rd = str(tmp_path / "rd")
    job = client.cloud_fit(_model(), rd, job_spec={"chief_config": CFG["CPU"], "worker_count": 1,
                                                  "worker_config": CFG["CPU"]},
                           x=x, y=y, epochs=5, batch_size=16, verbose=0)
    st = registry.wait(job, timeout=180)
    assert st["state"] == "SUCCEEDED", st
    assert os.path.isdir(os.path.join(rd, "output"))
    assert registry.get(job)["plan"]["world_size"] == 2
