"""The example scripts under examples/ run end to end on the CPU preset (ports of the reference's
core/tests/examples and core/tests/testdata programs; upstream only runs them in integration tests
against Google Cloud)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXAMPLES = os.path.join(ROOT, "examples")


def _run(args, cwd, tmp_path, timeout=600, **env_extra):
    env = dict(os.environ, PYTHONPATH=ROOT, TFC_JOB_ROOT=str(tmp_path / "jobs"),
               TFC_VIZIER_ROOT=str(tmp_path / "vizier"), CUDA_VISIBLE_DEVICES="", TFC_EXAMPLE_MACHINE="CPU",
               CHECKPOINT_DIR=str(tmp_path / "ckpt"), **env_extra)
    return subprocess.run([sys.executable] + args, cwd=cwd, env=env, capture_output=True, text=True,
                          timeout=timeout)


def test_keras_tuner_cifar_example(tmp_path):
    res = _run(["keras_tuner_cifar_example.py", "--path", str(tmp_path / "best"), "--max-trials", "2",
                "--samples", "128", "--directory", str(tmp_path / "tuner")], EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Search space summary" in res.stdout and "Tuner results summary" in res.stdout
    assert "Test accuracy:" in res.stdout and "Saving best model" in res.stdout
    assert os.path.exists(tmp_path / "best" / "config.json") and os.path.exists(tmp_path / "best" / "weights.h5")


def test_multi_file_example_travels_as_a_directory(tmp_path):
    res = _run(["scale_model.py"], os.path.join(EXAMPLES, "multi_file_example"), tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Job submitted successfully." in res.stdout
    assert "Epoch 2/2" in res.stdout and "eval:" in res.stdout          # remote(): 2 epochs, sibling module found


def test_notebook_entry_point(tmp_path):
    res = _run(["call_run_on_notebook.py"], EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Job submitted successfully." in res.stdout and "Epoch 2/2" in res.stdout
    assert "pip install" not in res.stdout


@pytest.mark.parametrize("script", ["mnist_example_using_fit.py", "mnist_example_using_ctl.py", "save_and_load.py"])
def test_plain_example_scripts(script, tmp_path):
    res = _run([script], EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
