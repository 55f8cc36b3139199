"""Failure detection (the reference delegates training-time fault tolerance
to AI Platform): a rank that dies mid-training (``TFC_FAULT=<rank>:<step>``) fails the job, the
healthy ranks are killed instead of hanging in the next collective, and the registry names the
failing rank."""
import os
import textwrap

from cloud_b200.core import containerize, deploy, machine_config
from cloud_b200.launch import registry

CFG = machine_config.COMMON_MACHINE_CONFIGS


def test_rank_crash_mid_training_fails_the_job(tmp_path, monkeypatch):
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    (tmp_path / "train.py").write_text(textwrap.dedent('''
        import numpy as np
        import cloud_b200 as tfc
        keras = tfc.keras
        tfc.distribute.experimental_set_strategy(tfc.distribute.MultiWorkerMirroredStrategy())
        m = keras.Sequential([keras.layers.Dense(4, activation="relu", input_shape=(8,)),
                              keras.layers.Dense(2, activation="softmax")])
        m.compile(optimizer="adam", loss="sparse_categorical_crossentropy")
        x = np.random.rand(256, 8).astype("float32"); y = np.random.randint(0, 2, 256)
        m.fit(x, y, batch_size=16, epochs=50, verbose=0)
        print("finished")
    '''))
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        uri = containerize.LocalContainerBuilder("train.py", None, CFG["CPU"], CFG["CPU"],
                                                 "local/p", "p").get_docker_image()
    finally:
        os.chdir(cwd)
    job = deploy.deploy_job("r", uri, CFG["CPU"], 1, CFG["CPU"], None, False,
                            env={"TFC_FAULT": "1:3"})
    st = registry.wait(job, timeout=180)
    assert st["state"] == "FAILED", st
    assert st["failed_rank"] == 1 and st["exit_code"] == 17
    assert st["seconds"] < 120
    logs = open(os.path.join(registry.job_dir(job), "logs", "rank0.log")).read()
    assert "finished" not in logs


TRAIN_WITH_BACKUP = '''
    import os
    import numpy as np
    import torch
    import cloud_b200 as tfc
    keras = tfc.keras
    tfc.distribute.experimental_set_strategy(tfc.distribute.MultiWorkerMirroredStrategy())
    torch.manual_seed(0)
    m = keras.Sequential([keras.layers.Dense(8, activation="relu", input_shape=(8,)),
                          keras.layers.Dense(2, activation="softmax")])
    m.compile(optimizer=keras.optimizers.Adam(1e-2), loss="sparse_categorical_crossentropy")
    rng = np.random.RandomState(0)
    x = rng.rand(128, 8).astype("float32"); y = (x[:, 0] > 0.5).astype("int64")
    out = os.environ["TRAIN_OUT"]
    cbs = [keras.callbacks.experimental.BackupAndRestore(os.path.join(out, "backup"))] \\
        if os.environ.get("USE_BACKUP") == "1" else []
    h = m.fit(x, y, batch_size=32, epochs=5, shuffle=False, verbose=2, callbacks=cbs)
    if m._is_chief:
        np.save(os.path.join(out, "weights.npy"), np.concatenate([w.reshape(-1) for w in m.get_weights()]))
        print("finished after attempt", os.environ.get("TFC_RESTART_COUNT"), "epochs run:", len(h.history["loss"]))
'''


def _run_training(tmp_path, name, env):
    import numpy as np
    work = tmp_path / name
    work.mkdir()
    (work / "train.py").write_text(textwrap.dedent(TRAIN_WITH_BACKUP))
    cwd = os.getcwd()
    os.chdir(work)
    try:
        uri = containerize.LocalContainerBuilder("train.py", None, CFG["CPU"], CFG["CPU"],
                                                 "local/p", "p").get_docker_image()
    finally:
        os.chdir(cwd)
    out = work / "out"
    out.mkdir()
    job = deploy.deploy_job("r", uri, CFG["CPU"], 1, CFG["CPU"], None, False,
                            env=dict(env, TRAIN_OUT=str(out)))
    st = registry.wait(job, timeout=240)
    log = open(os.path.join(registry.job_dir(job), "logs", "rank0.log")).read()
    weights = np.load(out / "weights.npy") if (out / "weights.npy").exists() else None
    return st, log, weights, out


def test_a_restarted_job_resumes_from_its_backup_and_ends_where_an_undisturbed_run_ends(tmp_path, monkeypatch):
    """TFC_JOB_RESTARTS + BackupAndRestore: rank 1 dies in the third epoch (step 9 of 4 per epoch),
    the supervisor starts the group again, training continues after epoch 2 and finishes with the
    weights of a run that was never interrupted."""
    import numpy as np
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    st, log, w_plain, _ = _run_training(tmp_path, "plain", {})
    assert st["state"] == "SUCCEEDED" and st["restarts"] == 0, st
    st, log, w_resumed, out = _run_training(tmp_path, "faulty", {"TFC_FAULT": "1:9", "USE_BACKUP": "1",
                                                               "TFC_JOB_RESTARTS": "2"})
    assert st["state"] == "SUCCEEDED" and st["restarts"] == 1, st
    assert st["failures"][0]["rank"] == 1 and st["failures"][0]["exit_code"] == 17
    assert "---- restart 1 of 2 ----" in log
    assert "BackupAndRestore: resuming after epoch 2" in log
    assert "finished after attempt 1 epochs run: 3" in log              # epochs 3, 4, 5 only
    np.testing.assert_allclose(w_resumed, w_plain, rtol=1e-5, atol=1e-6)
    assert not (out / "backup").exists()                                 # removed after a clean end


def test_without_restarts_left_the_job_fails_and_keeps_its_backup(tmp_path, monkeypatch):
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    st, log, weights, out = _run_training(tmp_path, "norestart", {"TFC_FAULT": "0:6", "USE_BACKUP": "1"})
    assert st["state"] == "FAILED" and st["restarts"] == 0 and weights is None
    assert (out / "backup" / "backup.json").exists()                     # a later run can pick it up
