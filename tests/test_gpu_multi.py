"""Multi-GPU tests (need >= 2 B200 on the box: `gpurun --gpus 2`): the peer-memory collectives
against NCCL and data-parallel training against a single-replica oracle."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(script, nproc, args=(), env_extra=None, port=29700, timeout=600):
    env = dict(os.environ)
    env.update(env_extra or {})
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), script, *args]
    return subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_p2p_collectives_match_nccl(tmp_path):
    n = min(_ngpu(), 8)
    res = _torchrun(os.path.join(ROOT, "bench", "allreduce_sweep.py"), n,
                    args=["--check-only", "--out", str(tmp_path / "sweep.json")], port=29701)
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert "CHECKS_OK" in res.stdout


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("eager", ["1", "0"])
def test_mirrored_training_matches_single_gpu(tmp_path, eager):
    n = min(_ngpu(), 8)
    res = _torchrun(os.path.join(ROOT, "tests", "dist_train_check.py"), n,
                    env_extra={"CHECK_OUT": str(tmp_path), "CHECK_EAGER": eager}, port=29702 + int(eager))
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert "backend=p2p" in res.stdout, res.stdout[-2000:]


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_custom_training_loop_matches_single_gpu():
    n = min(_ngpu(), 8)
    res = _torchrun(os.path.join(ROOT, "tests", "dist_ctl_check.py"), n, port=29706)
    assert res.returncode == 0, res.stdout[-4000:] + res.stderr[-4000:]
    assert "CTL OK" in res.stdout and "backend=p2p" in res.stdout, res.stdout[-2000:]
