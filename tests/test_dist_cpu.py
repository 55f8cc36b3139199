"""Multi-process CPU tests (world_size 2 and 3): the host-side logic of the data-parallel engine -
bucketing, hooks, broadcast, metric reduction, chief-only saving - without a GPU, over both CPU
transports: the shared-memory collectives that mirror the GPU kernels' schedules and flag protocol
(default for CPU ranks) and plain torch.distributed / gloo.  The reference
never really tests multi-node (mocked, TF_CONFIG-faked or process pools)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(script, nproc, env_extra, port, timeout=240):
    env = dict(os.environ)
    env.update(env_extra)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), script]
    return subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)


@pytest.mark.parametrize("nproc,backend", [(2, "shm"), (3, "shm"), (2, "dist"), (3, "dist")])
def test_mirrored_equals_single_replica_cpu(tmp_path, nproc, backend):
    env = {"CHECK_DEVICE": "cpu", "CUDA_VISIBLE_DEVICES": "", "CHECK_OUT": str(tmp_path),
           "CHECK_RATIO_METRIC": "1"}
    if backend != "shm":                       # shm is what CPU ranks pick by themselves
        env["TFC_COMM_BACKEND"] = backend
    res = _torchrun(os.path.join(ROOT, "tests", "dist_train_check.py"), nproc, env,
                    port=29610 + nproc + (10 if backend == "dist" else 0))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "OK world=%d backend=%s" % (nproc, backend) in res.stdout


def test_shared_memory_collectives_across_processes(tmp_path):
    res = _torchrun(os.path.join(ROOT, "tests", "shm_collectives_check.py"), 4,
                    {"CUDA_VISIBLE_DEVICES": ""}, port=29631)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "SHM OK world=4" in res.stdout


def test_a_rank_that_never_arrives_fails_the_others_instead_of_hanging(tmp_path):
    res = _torchrun(os.path.join(ROOT, "tests", "shm_collectives_check.py"), 3,
                    {"CUDA_VISIBLE_DEVICES": "", "SHM_CHECK_ABSENT_RANK": "2"}, port=29632)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert res.stdout.count("WATCHDOG OK") == 2 and "for rank 2" in res.stdout


def test_custom_training_loop_equals_single_replica_cpu(tmp_path):
    res = _torchrun(os.path.join(ROOT, "tests", "dist_ctl_check.py"), 2,
                    {"CHECK_DEVICE": "cpu", "CUDA_VISIBLE_DEVICES": ""}, port=29641)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "CTL OK world=2 backend=shm" in res.stdout
