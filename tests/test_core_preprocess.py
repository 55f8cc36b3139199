"""Port of core/tests/unit/preprocess_test.py: the generated wrapper is pinned line by line for
the OneDevice / Mirrored / MultiWorker / TPU cases and for notebooks (parsed as JSON here instead
of through a mocked nbconvert PythonExporter)."""
import os

import pytest

from cloud_b200.core import machine_config, preprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ENTRY = os.path.join(HERE, "testdata", "mnist_example_using_fit.py")
NOTEBOOK = os.path.join(HERE, "testdata", "mnist_example_using_fit.ipynb")
CFG = machine_config.COMMON_MACHINE_CONFIGS

HEADER = ["import os\n", "import cloud_b200 as tfc\n", 'os.environ["TF_KERAS_RUNNING_REMOTELY"]="1"\n']
SET = "tfc.distribute.experimental_set_strategy(strategy)\n"
EXEC = 'exec(open("mnist_example_using_fit.py").read())\n'


def _lines(path):
    with open(path) as f:
        out = f.readlines()
    os.remove(path)
    return out


def test_auto_one_device_strategy_even_for_cpu_preset():
    path = preprocess.get_preprocessed_entry_point(ENTRY, CFG["CPU"], CFG["CPU"], 0, "auto")
    assert _lines(path) == HEADER + [
        "strategy = tfc.distribute.OneDeviceStrategy(device='/gpu:0')\n", SET, EXEC]


def test_auto_mirrored_strategy():
    path = preprocess.get_preprocessed_entry_point(ENTRY, CFG["K80_4X"], CFG["K80_1X"], 0, "auto")
    assert _lines(path) == HEADER + ["strategy = tfc.distribute.MirroredStrategy()\n", SET, EXEC]


def test_auto_multi_worker_strategy():
    path = preprocess.get_preprocessed_entry_point(ENTRY, CFG["K80_1X"], CFG["K80_1X"], 2, "auto")
    assert _lines(path) == HEADER + [
        "strategy = tfc.distribute.experimental.MultiWorkerMirroredStrategy()\n", SET, EXEC]


def test_auto_tpu_worker_emits_resolver_and_tpu_strategy():
    path = preprocess.get_preprocessed_entry_point(ENTRY, CFG["CPU"], CFG["TPU"], 1, "auto")
    lines = _lines(path)
    assert lines[:3] == HEADER
    assert lines[3] == "def wait_for_tpu_cluster_resolver_ready():\n"
    assert "resolver = wait_for_tpu_cluster_resolver_ready()\n" in lines
    assert "strategy = tfc.distribute.experimental.TPUStrategy(resolver)\n" in lines
    assert lines[-2:] == [SET, EXEC]


def test_no_strategy_injection_when_strategy_is_none():
    path = preprocess.get_preprocessed_entry_point(ENTRY, CFG["K80_4X"], CFG["K80_1X"], 0, None)
    assert _lines(path) == HEADER + [EXEC]


def test_entry_point_none_uses_calling_script(monkeypatch):
    monkeypatch.setattr("sys.argv", ["/somewhere/caller_script.py"])
    path = preprocess.get_preprocessed_entry_point(None, CFG["CPU"], CFG["CPU"], 0, "auto")
    assert _lines(path)[-1] == 'exec(open("caller_script.py").read())\n'


def test_notebook_entry_point_inlines_code_cells_without_magics():
    path = preprocess.get_preprocessed_entry_point(NOTEBOOK, CFG["CPU"], CFG["CPU"], 0, "auto")
    assert _lines(path) == HEADER + [
        "strategy = tfc.distribute.OneDeviceStrategy(device='/gpu:0')\n", SET,
        "import numpy as np\n", "print('from notebook', np.__name__)\n", "x = 1\n", "print(x)\n"]


def test_unreadable_notebook_raises_runtime_error(tmp_path):
    bad = tmp_path / "broken.ipynb"
    bad.write_text("{not json")
    with pytest.raises(RuntimeError, match="Unable to access iPython notebook"):
        preprocess.get_preprocessed_entry_point(str(bad), CFG["CPU"], CFG["CPU"], 0, "auto")


def test_generated_wrapper_runs_and_sets_global_strategy(tmp_path, monkeypatch):
    script = tmp_path / "user.py"
    script.write_text("import cloud_b200 as tfc\nprint('S=', type(tfc.distribute.get_strategy()).__name__, tfc.remote())\n")
    path = preprocess.get_preprocessed_entry_point(str(script), CFG["CPU"], CFG["CPU"], 0, "auto")
    import subprocess
    import sys
    root = os.path.dirname(HERE)
    env = dict(os.environ, PYTHONPATH=root, CUDA_VISIBLE_DEVICES="")
    env.pop("TF_KERAS_RUNNING_REMOTELY", None)
    out = subprocess.run([sys.executable, path], cwd=tmp_path, env=env, capture_output=True, text=True)
    os.remove(path)
    assert out.returncode == 0, out.stderr
    assert "S= OneDeviceStrategy True" in out.stdout
