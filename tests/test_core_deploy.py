"""Port of core/tests/unit/deploy_test.py and integration/run_on_script_test.py: the rank plan that
replaces the CAIP request body, the printed messages, failure paths, and `tfc.run` end to end on
CPU ranks (only sys.exit mocked, like upstream :37,50-137)."""
import json
import os
import textwrap
import time
from unittest import mock

import pytest

import cloud_b200 as tfc
from cloud_b200.core import containerize, deploy, machine_config
from cloud_b200.launch import plan as plan_mod
from cloud_b200.launch import registry

CFG = machine_config.COMMON_MACHINE_CONFIGS
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(autouse=True)
def _env(monkeypatch, tmp_path):
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    monkeypatch.setenv("TFC_PROJECT", "my-project")
    monkeypatch.delenv("TF_KERAS_RUNNING_REMOTELY", raising=False)


def test_plan_chief_and_workers():
    p = plan_mod.plan_ranks(CFG["K80_4X"], 2, CFG["K80_1X"], visible_gpus=8)
    assert p["world_size"] == 6 and p["gpu_ids"] == [0, 1, 2, 3, 4, 5]
    assert [r["task_type"] for r in p["ranks"]] == ["chief"] * 4 + ["worker", "worker"]
    assert [r["task_index"] for r in p["ranks"]] == [0, 0, 0, 0, 0, 1]
    assert [r["rank_in_task"] for r in p["ranks"]] == [0, 1, 2, 3, 0, 0]
    cfg = json.loads(p["ranks"][5]["tf_config"])
    assert cfg["task"] == {"type": "worker", "index": 1}
    assert set(cfg["cluster"]) == {"chief", "worker"} and len(cfg["cluster"]["worker"]) == 2


def test_plan_cpu_replicas_get_one_rank_each_and_no_gpu():
    p = plan_mod.plan_ranks(CFG["CPU"], 1, CFG["CPU"], visible_gpus=0)
    assert p["world_size"] == 2 and not p["uses_gpu"] and [r["gpu"] for r in p["ranks"]] == [None, None]


def test_plan_errors():
    with pytest.raises(plan_mod.RankPlanError, match=r"needs 8 GPUs .* only 4 are visible"):
        plan_mod.plan_ranks(CFG["K80_8X"], 0, None, visible_gpus=4)
    with pytest.raises(plan_mod.RankPlanError, match=r"no TPU on this box"):
        plan_mod.plan_ranks(CFG["CPU"], 1, CFG["TPU"], visible_gpus=8)


def _stage(tmp_path, body="print('hello from', __import__('os').environ['RANK'])\n"):
    (tmp_path / "train.py").write_text(body)
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        b = containerize.LocalContainerBuilder("train.py", None, CFG["CPU"], CFG["CPU"],
                                               "local/my-project", "my-project")
        uri = b.get_docker_image()
    finally:
        os.chdir(cwd)
    return uri


def test_request_dict_keeps_caip_shape(tmp_path, monkeypatch):
    monkeypatch.setenv("TFC_FAKE_GPUS", "8")
    uri = _stage(tmp_path)
    req = deploy._create_request_dict("tf_cloud_train_x", "us-central1", uri, CFG["K80_4X"], 2,
                                      CFG["K80_1X"], ["1000"], job_labels={"team": "a"})
    ti = req["trainingInput"]
    assert req["jobId"] == "tf_cloud_train_x" and req["labels"] == {"team": "a"}
    assert ti["scaleTier"] == "custom" and ti["masterType"] == "n1-standard-16"
    assert ti["masterConfig"] == {"imageUri": uri, "acceleratorConfig": {"count": "4", "type": "NVIDIA_TESLA_K80"}}
    assert ti["workerCount"] == "2" and ti["workerType"] == "n1-standard-8"
    assert ti["args"] == ["1000"] and ti["use_chief_in_tf_config"] is True
    assert req["plan"]["world_size"] == 6
    no_workers = deploy._create_request_dict("j", "r", uri, CFG["K80_1X"], 0, CFG["K80_1X"], None, {})
    assert "workerType" not in no_workers["trainingInput"] and "args" not in no_workers["trainingInput"]
    assert "labels" not in no_workers


def test_deploy_job_prints_and_runs(tmp_path, capsys):
    uri = _stage(tmp_path)
    job_id = deploy.deploy_job("us-central1", uri, CFG["CPU"], 1, CFG["CPU"], None, True)
    out = capsys.readouterr().out.splitlines()
    assert job_id.startswith("tf_cloud_train_")
    assert out[0] == "Job submitted successfully."
    assert out[1] == "Your job ID is:  " + job_id
    assert out[2] == "Please access your job logs at the following URL:"
    assert out[3].startswith("file://") and out[3].endswith(os.path.join(job_id, "logs"))
    assert "Streaming job logs: " in out
    assert "[rank0] hello from 0" in out and "[rank1] hello from 1" in out
    assert registry.status(job_id)["state"] == "SUCCEEDED"


def test_failing_rank_fails_job_and_is_reported(tmp_path, capsys):
    uri = _stage(tmp_path, "import os, sys, time\nif os.environ['RANK'] == '1':\n    print('boom'); sys.exit(3)\ntime.sleep(30)\n")
    job_id = deploy.deploy_job("us-central1", uri, CFG["CPU"], 1, CFG["CPU"], None, True)
    st = registry.status(job_id)
    assert st["state"] == "FAILED" and st["failed_rank"] == 1 and st["exit_code"] == 3
    assert "boom" in st["error_tail"]
    assert st["seconds"] < 20                      # the healthy rank was killed, not waited for
    assert "Job failed (rank 1, exit code 3)." in capsys.readouterr().out


def test_stalled_rank_is_detected_by_its_heartbeat(tmp_path, capsys, monkeypatch):
    """A rank that beat once and then hangs (a wedged collective) fails the job (section 5.3)."""
    monkeypatch.setenv("TFC_HEARTBEAT_TIMEOUT", "1.5")
    script = ("import os, time\n"
              "hb = os.path.join(os.environ['TFC_JOB_DIR'], 'heartbeat', 'rank' + os.environ['RANK'])\n"
              "os.makedirs(os.path.dirname(hb), exist_ok=True)\n"
              "for i in range(100):\n"
              "    if os.environ['RANK'] == '0' or i == 0:\n"
              "        open(hb, 'a').close(); os.utime(hb, None)\n"
              "    time.sleep(0.3)\n")
    uri = _stage(tmp_path, script)
    job_id = deploy.deploy_job("us-central1", uri, CFG["CPU"], 1, CFG["CPU"], None, True)
    st = registry.status(job_id)
    assert st["state"] == "FAILED" and st["failed_rank"] == 1
    assert "no heartbeat" in str(st["exit_code"])
    assert st["seconds"] < 20


def test_submission_error_is_reported(tmp_path, capsys):
    uri = _stage(tmp_path)
    with mock.patch.object(deploy, "_start_supervisor", side_effect=OSError("no fork")):
        with pytest.raises(OSError):
            deploy.deploy_job("us-central1", uri, CFG["CPU"], 0, CFG["CPU"], None, False)
    assert "There was an error submitting the job." in capsys.readouterr().out


def test_stream_logs_error_is_reported(tmp_path, capsys):
    uri = _stage(tmp_path)
    with mock.patch.object(registry, "log_paths", side_effect=OSError("gone")):
        with pytest.raises(OSError):
            deploy.deploy_job("us-central1", uri, CFG["CPU"], 0, CFG["CPU"], None, True)
    assert "There was an error streaming the job logs." in capsys.readouterr().out


def test_job_id_format():
    jid = deploy._generate_job_id()
    assert jid.startswith("tf_cloud_train_") and "-" not in jid


# ---- run() -------------------------------------------------------------------------------------
def _write_script(tmp_path):
    (tmp_path / "train.py").write_text(textwrap.dedent('''
        import os
        import cloud_b200 as tfc
        st = tfc.distribute.get_strategy()
        print("RUN", type(st).__name__, st.num_replicas_in_sync, tfc.remote(), os.environ.get("MARK", ""))
        tfc.run()    # nested call inside the job is a no-op (run.py:134-135)
    '''))


@pytest.mark.parametrize("kwargs,expect", [
    (dict(chief_config=CFG["CPU"]), "RUN OneDeviceStrategy 1 True"),
    (dict(chief_config=CFG["CPU"], worker_config=CFG["CPU"], worker_count=1),
     "RUN MultiWorkerMirroredStrategy 2 True"),
    (dict(chief_config=CFG["CPU"], distribution_strategy=None), "RUN OneDeviceStrategy 1 True"),
    (dict(chief_config=CFG["CPU"], docker_image_bucket_name="bkt", job_labels={"job": "on_script"}),
     "RUN OneDeviceStrategy 1 True"),
])
def test_run_on_script(tmp_path, monkeypatch, capsys, kwargs, expect):
    _write_script(tmp_path)
    monkeypatch.chdir(tmp_path)
    with mock.patch("sys.exit") as exit_mock:
        tfc.run(entry_point="train.py", stream_logs=True, **kwargs)
    exit_mock.assert_called_once_with(0)
    out = capsys.readouterr().out
    assert expect in out, out
    jobs = registry.list_jobs()
    assert len(jobs) == 1 and registry.status(jobs[0])["state"] == "SUCCEEDED"


def test_run_rejects_unknown_kwargs_and_is_noop_when_remote(monkeypatch):
    with pytest.raises(TypeError, match=r"Unknown keyword arguments"):
        tfc.run(entry_point="x.py", foo=1)
    monkeypatch.setenv("TF_KERAS_RUNNING_REMOTELY", "1")
    assert tfc.remote() is True
    assert tfc.run(entry_point="does_not_exist.py") is None


def test_run_with_tpu_worker_reports_no_tpu(tmp_path, monkeypatch):
    _write_script(tmp_path)
    monkeypatch.chdir(tmp_path)
    with pytest.raises(plan_mod.RankPlanError, match=r"no TPU on this box"):
        tfc.run(entry_point="train.py", chief_config=CFG["CPU"], worker_config=CFG["TPU"],
                worker_count=1, docker_base_image="custom")


def test_job_cli_lists_describes_streams_and_cancels(tmp_path, capsys):
    """`python -m cloud_b200.launch ...`: the local stand-in for `gcloud ai-platform jobs ...`."""
    from cloud_b200.launch import __main__ as cli
    uri = _stage(tmp_path, "print('hello from the job')\n")
    done = deploy.deploy_job("us-central1", uri, CFG["CPU"], 0, CFG["CPU"], None, False)
    assert registry.wait(done, timeout=60)["state"] == "SUCCEEDED"
    slow_uri = _stage(tmp_path, "import time\nprint('started', flush=True)\ntime.sleep(120)\n")
    slow = deploy.deploy_job("us-central1", slow_uri, CFG["CPU"], 0, CFG["CPU"], None, False)
    t0 = time.time()
    while registry.status(slow).get("state") != "RUNNING" and time.time() - t0 < 30:
        time.sleep(0.1)
    capsys.readouterr()
    assert cli.main(["list"]) == 0
    out = capsys.readouterr().out
    assert done in out and "SUCCEEDED" in out and slow in out and "RUNNING" in out
    assert cli.main(["describe", done]) == 0
    desc = json.loads(capsys.readouterr().out)
    assert desc["status"]["state"] == "SUCCEEDED" and desc["plan"]["world_size"] == 1 and desc["logs"]
    assert cli.main(["stream-logs", done]) == 0
    assert "hello from the job" in capsys.readouterr().out
    assert cli.main(["cancel", slow]) == 0
    assert registry.status(slow)["state"] == "CANCELLED"
    assert cli.main(["cancel", slow]) == 0 and "already CANCELLED" in capsys.readouterr().out
    assert cli.main(["describe", "tf_cloud_train_nope"]) == 2
    assert cli.main(["topology"]) == 0
