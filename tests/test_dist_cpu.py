"""Multi-process CPU tests (gloo, world_size 2): the host-side logic of the data-parallel engine -
bucketing, hooks, broadcast, metric reduction, chief-only saving - without a GPU.  The reference
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


@pytest.mark.parametrize("nproc", [2, 3])
def test_mirrored_equals_single_replica_cpu(tmp_path, nproc):
    res = _torchrun(os.path.join(ROOT, "tests", "dist_train_check.py"), nproc,
                    {"CHECK_DEVICE": "cpu", "CUDA_VISIBLE_DEVICES": "", "CHECK_OUT": str(tmp_path)},
                    port=29610 + nproc)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "OK world=%d backend=dist" % nproc in res.stdout
