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
