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


# -- the reference's remaining example programs (core/tests/examples/*.py) ------------------------
def test_ctl_script_with_a_user_managed_strategy_and_one_worker(tmp_path):
    res = _run(["call_run_on_script_with_keras_ctl.py"], EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Job submitted successfully." in res.stdout and "Epoch 1: loss" in res.stdout


def test_save_and_load_script_gets_its_arguments(tmp_path):
    out = tmp_path / "saved"
    res = _run(["call_run_on_script_with_keras_save_and_load.py", "--path", str(out)], EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "save_and_load OK" in res.stdout and os.path.isdir(out)


def test_staged_image_build_when_a_bucket_is_named(tmp_path):
    res = _run(["call_run_on_script_with_keras_fit_cloud_build.py", "--bucket_name", "staging-a"],
               EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Epoch 2/2" in res.stdout
    assert os.path.isdir(tmp_path / "jobs" / "buckets" / "staging-a") or "staging-a" in res.stdout + res.stderr


def test_tuner_search_as_one_job(tmp_path):
    res = _run(["call_run_on_script_with_keras_tuner_search.py", "--path", str(tmp_path / "best"),
                "--max-trials", "2", "--samples", "128"], EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Saving best model" in res.stdout and os.path.exists(tmp_path / "best" / "weights.h5")


def test_model_search_inside_a_job(tmp_path):
    res = _run(["call_run_within_script_with_autokeras.py", "--path", str(tmp_path / "ak"), "--samples", "256",
                "--epochs", "2"], EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Accuracy:" in res.stdout and os.path.exists(tmp_path / "ak" / "model.h5")


def test_hyperband_and_bayesian_tuners_example(tmp_path):
    res = _run(["keras_tuner_hyperband_example.py", "--samples", "384", "--max-epochs", "4", "--bayesian-trials", "3",
                "--directory", str(tmp_path / "tuners")], EXAMPLES, tmp_path)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Hyperband ran 10 trials" in res.stdout and "Bayesian optimisation: best validation accuracy" in res.stdout


# -- notebooks, run headless by cloud_b200.utils.notebook ----------------------------------------
def _run_notebook(name, tmp_path, **env):
    return _run(["-m", "cloud_b200.utils.notebook", name], EXAMPLES, tmp_path, TFC_EXAMPLE_EPOCHS="2", **env)


def test_notebook_that_submits_itself(tmp_path):
    res = _run_notebook("call_run_within_nb_on_colab.ipynb", tmp_path, TFC_SYNTHETIC_SAMPLES="512",
                        TFC_EXAMPLE_MODEL_PATH=str(tmp_path / "colab_model"))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    # the local look (1 epoch), the job (2 epochs, learning rate printed by the callback), then the
    # notebook goes on after the submission and evaluates what the job saved
    assert "Job submitted successfully." in res.stdout and "[rank0] Epoch 2/2" in res.stdout
    assert "Learning rate for epoch 1 is 0.001" in res.stdout
    assert os.path.exists(tmp_path / "colab_model" / "weights.h5")
    assert res.stdout.rstrip().splitlines()[-1].startswith("loss:")


def test_transfer_learning_notebook_with_a_frozen_base(tmp_path):
    res = _run_notebook("dogs_classification.ipynb", tmp_path, TFC_SYNTHETIC_SAMPLES="128",
                        TFC_EXAMPLE_IMG_SIZE="32", TFC_EXAMPLE_LOCAL_BATCHES="1",
                        TFC_EXAMPLE_MODEL_PATH=str(tmp_path / "dogs"))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "Number of classes: 120" in res.stdout and "[rank0] Epoch 2/2" in res.stdout
    assert os.path.exists(tmp_path / "dogs" / "final" / "weights.h5")
    assert os.path.isdir(tmp_path / "dogs" / "save_at_2")           # ModelCheckpoint per epoch


def test_cloud_fit_notebook(tmp_path):
    res = _run_notebook("cloud_fit.ipynb", tmp_path, TFC_SYNTHETIC_SAMPLES="128",
                        TFC_EXAMPLE_REMOTE_DIR=str(tmp_path / "remote"))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert res.stdout.count("SUCCEEDED") >= 2


def test_cloud_tuner_notebook(tmp_path):
    res = _run_notebook("ai_platform_optimizer_tuner.ipynb", tmp_path, TFC_SYNTHETIC_SAMPLES="256",
                        TFC_EXAMPLE_MAX_TRIALS="2", TFC_EXAMPLE_TUNER_DIR=str(tmp_path / "tuner_nb"))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert res.stdout.count("Search space summary") == 2 and res.stdout.count("Results summary") >= 4
    assert "tuner3 exit code 0" in res.stdout
