"""GPU leases: concurrent jobs on one box never share a GPU, jobs queue (FIFO) when the box is full,
dead holders do not keep their GPUs, queued jobs can be cancelled.  New tier: on Google Cloud every
job of the reference had its own machines (core/deploy.py:82-88), so it has nothing like this."""
import os
import subprocess
import sys
import time

import pytest

from cloud_b200.core import containerize, deploy, machine_config
from cloud_b200.launch import leases, registry
from cloud_b200.launch import __main__ as cli

CFG = machine_config.COMMON_MACHINE_CONFIGS


@pytest.fixture(autouse=True)
def _env(monkeypatch, tmp_path):
    monkeypatch.setenv("TFC_JOB_ROOT", str(tmp_path / "jobs"))
    monkeypatch.setenv("TFC_PROJECT", "my-project")
    monkeypatch.setenv("TFC_FAKE_GPUS", "2")           # the box "has" two GPUs (ranks run on the CPU)
    monkeypatch.delenv("TF_KERAS_RUNNING_REMOTELY", raising=False)


def test_lease_table_basics():
    assert leases.try_acquire("a", 1, 2) == [0]
    assert leases.try_acquire("b", 1, 2) == [1]
    assert leases.try_acquire("c", 1, 2) is None                     # full: c waits
    assert leases.try_acquire("d", 1, 2) is None
    snap = leases.snapshot(2)
    assert snap["free"] == [] and [q["owner"] for q in snap["queue"]] == ["c", "d"]
    assert leases.release("a") == [0]
    assert leases.try_acquire("d", 1, 2) is None                     # FIFO: c asked first
    assert leases.try_acquire("c", 1, 2) == [0]
    leases.release("b")
    assert leases.try_acquire("d", 1, 2) == [1]
    with pytest.raises(ValueError):
        leases.try_acquire("e", 3, 2)
    assert leases.try_acquire("nothing", 0, 2) == []
    leases.cancel("d")
    leases.release("c"), leases.release("d")
    assert leases.snapshot(2)["free"] == [0, 1]


def test_a_dead_holder_loses_its_lease():
    child = subprocess.Popen([sys.executable, "-c", "import time; time.sleep(60)"])
    assert leases.try_acquire("ghost", 2, 2, pid=child.pid) == [0, 1]
    assert leases.try_acquire("next", 1, 2) is None
    child.kill()
    child.wait()
    assert leases.try_acquire("next", 1, 2) == [0]                   # reclaimed without any clean-up call
    with pytest.raises(leases.LeaseTimeout):
        leases.acquire("greedy", 2, 2, timeout=0.5, poll=0.05)
    assert leases.snapshot(2)["queue"] == []                         # a timed-out request leaves the queue


def _stage(tmp_path, body):
    (tmp_path / "train.py").write_text(body)
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        return containerize.LocalContainerBuilder("train.py", None, CFG["B200_1X"], CFG["B200_1X"],
                                                  "local/my-project", "my-project").get_docker_image()
    finally:
        os.chdir(cwd)


BODY = ("import os, time\n"
        "print('gpu', os.environ['CUDA_VISIBLE_DEVICES'], 'from', time.time(), flush=True)\n"
        "time.sleep(2.0)\n"
        "print('until', time.time(), flush=True)\n")


def _interval(job_id):
    text = open(registry.log_paths(job_id)[0]).read().split()
    return text[1], float(text[3]), float(text[5])


def test_three_one_gpu_jobs_on_a_two_gpu_box(tmp_path, capsys):
    uri = _stage(tmp_path, BODY)
    jobs = [deploy.deploy_job("us-central1", uri, CFG["B200_1X"], 0, CFG["B200_1X"], None, False)
            for _ in range(3)]
    seen_queued = False
    deadline = time.time() + 60
    while time.time() < deadline:
        states = [registry.status(j) for j in jobs]
        seen_queued |= any(s.get("state") == "QUEUED" and s.get("waiting_for_gpus") == 1 for s in states)
        if all(s.get("state") in registry.TERMINAL for s in states):
            break
        time.sleep(0.1)
    assert [s["state"] for s in states] == ["SUCCEEDED"] * 3, states
    assert seen_queued, "the third job never waited for a GPU"
    runs = [_interval(j) for j in jobs]
    assert {g for g, _a, _b in runs} == {"0", "1"}
    for i in range(3):
        assert states[i]["gpus"] == [int(runs[i][0])]
        for k in range(i + 1, 3):
            same_gpu = runs[i][0] == runs[k][0]
            overlap = runs[i][1] < runs[k][2] and runs[k][1] < runs[i][2]
            assert not (same_gpu and overlap), "two jobs used GPU {} at the same time".format(runs[i][0])
    assert leases.snapshot(2)["free"] == [0, 1]
    assert cli.main(["gpus"]) == 0
    assert "GPU" in capsys.readouterr().out


def test_a_two_gpu_job_waits_for_both_and_a_queued_job_can_be_cancelled(tmp_path, capsys):
    uri = _stage(tmp_path, BODY)
    first = deploy.deploy_job("us-central1", uri, CFG["B200_1X"], 0, CFG["B200_1X"], None, False)
    time.sleep(0.5)
    both = deploy.deploy_job("us-central1", uri, CFG["B200_2X"], 0, CFG["B200_1X"], None, False)
    deadline = time.time() + 20
    while time.time() < deadline and registry.status(both).get("waiting_for_gpus") != 2:
        time.sleep(0.05)
    assert registry.status(both).get("state") == "QUEUED"
    assert cli.main(["gpus"]) == 0
    assert "waiting: {} wants 2 GPU(s)".format(both) in capsys.readouterr().out
    third = deploy.deploy_job("us-central1", uri, CFG["B200_1X"], 0, CFG["B200_1X"], None, False)
    time.sleep(0.5)
    assert registry.status(third).get("state") == "QUEUED"            # behind the two-GPU job, although a GPU is free
    assert cli.main(["cancel", third]) == 0
    assert registry.status(third)["state"] == "CANCELLED"
    assert registry.wait(both, timeout=60)["state"] == "SUCCEEDED"
    assert registry.status(both)["gpus"] == [0, 1]
    a, b = _interval(first), _interval(both)
    assert b[1] >= a[2] - 0.05                                        # started only after the first finished
    assert b[0] == "0,1"


def test_queue_timeout_fails_the_job(tmp_path, monkeypatch):
    uri = _stage(tmp_path, BODY.replace("2.0", "4.0"))
    hog = deploy.deploy_job("us-central1", uri, CFG["B200_2X"], 0, CFG["B200_1X"], None, False)
    time.sleep(0.5)
    monkeypatch.setenv("TFC_QUEUE_TIMEOUT", "1")
    late = deploy.deploy_job("us-central1", uri, CFG["B200_1X"], 0, CFG["B200_1X"], None, False)
    st = registry.wait(late, timeout=30)
    assert st["state"] == "FAILED" and "waited" in st["error_tail"] and "free GPU" in st["error_tail"]
    assert registry.wait(hog, timeout=60)["state"] == "SUCCEEDED"


def _report_gpu(tuner_id):
    import os
    import time
    time.sleep(0.3)
    return {"tuner": tuner_id, "gpus": os.environ["CUDA_VISIBLE_DEVICES"]}


def test_trial_scheduler_shares_the_box_with_a_running_job(tmp_path):
    from cloud_b200.tuner import TrialScheduler
    holder = subprocess.Popen([sys.executable, "-c", "import time; time.sleep(60)"])
    try:
        assert leases.try_acquire("some_job", 1, 2, pid=holder.pid) == [0]
        records = TrialScheduler(gpus_per_trial=1).run(_report_gpu, num_workers=3, log_dir=str(tmp_path / "logs"))
        assert [r["exit_code"] for r in records] == [0, 0, 0]
        assert {tuple(r["gpus"]) for r in records} == {(1,)}            # GPU 0 belongs to the job
        assert leases.snapshot(2)["gpus"].keys() == {"0"}               # the scheduler gave its slot back
    finally:
        holder.kill()
        holder.wait()
    records = TrialScheduler(gpus_per_trial=1).run(_report_gpu, num_workers=4, log_dir=str(tmp_path / "logs2"))
    assert {tuple(r["gpus"]) for r in records} == {(0,), (1,)}          # the whole box again


def _lease_worker(args):
    root, worker, rounds = args
    import os
    import time
    os.environ["TFC_JOB_ROOT"] = root
    from cloud_b200.launch import leases as ls
    events = []
    for i in range(rounds):
        owner = "w{}_{}".format(worker, i)
        got = ls.acquire(owner, 1 + (worker % 2), 3, timeout=60, poll=0.01)
        t0 = time.time()
        time.sleep(0.03)
        t1 = time.time()
        ls.release(owner)
        events.append((got, t0, t1))
    return events


def test_lease_table_survives_concurrent_processes(tmp_path):
    """Six processes fight over three GPUs (some want two at once): no GPU is ever held twice at the
    same time and everybody gets served (the table is one JSON file under flock)."""
    import multiprocessing as mp
    root = str(tmp_path / "jobs")
    ctx = mp.get_context("spawn")
    with ctx.Pool(6) as pool:
        results = pool.map(_lease_worker, [(root, w, 4) for w in range(6)])
    holds = [(g, t0, t1) for events in results for got, t0, t1 in events for g in got]
    assert len(holds) == 4 * (3 * 1 + 3 * 2)
    for gpu in range(3):
        mine = sorted((t0, t1) for g, t0, t1 in holds if g == gpu)
        assert mine, "GPU {} was never handed out".format(gpu)
        for (a0, a1), (b0, _b1) in zip(mine, mine[1:]):
            assert a1 <= b0 + 1e-3, "GPU {} was held twice at once".format(gpu)
    assert leases.snapshot(3)["free"] == [0, 1, 2] and leases.snapshot(3)["queue"] == []
