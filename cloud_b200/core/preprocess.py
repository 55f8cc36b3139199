"""Strategy injection and entry-point / notebook wrapping
(reference: src/python/tensorflow_cloud/core/preprocess.py:38-262).

The wrapper script has the reference's shape, line for line - header, strategy, global
``experimental_set_strategy``, then the user code - but names ``cloud_b200`` strategies:

    import os
    import cloud_b200 as tfc
    os.environ["TF_KERAS_RUNNING_REMOTELY"]="1"
    strategy = tfc.distribute.MirroredStrategy()
    tfc.distribute.experimental_set_strategy(strategy)
    exec(open("mnist_example.py").read())

Selection rule (preprocess.py:124-147): workers > 0 -> (TPU resolver +) TPUStrategy for a TPU
worker, else MultiWorkerMirroredStrategy; chief.accelerator_count > 1 -> MirroredStrategy; else
OneDeviceStrategy('/gpu:0') - even for the CPU preset, like the reference (the strategy maps
'/gpu:0' to the CPU when the replica has no GPU).  Notebooks are read as JSON (nbconvert is not
installed here): code cells are concatenated and lines starting with ``!``, ``%`` or ``#`` dropped.
"""
from __future__ import annotations

import json
import os
import sys
import tempfile

from . import machine_config

FRAMEWORK_IMPORT = "import cloud_b200 as tfc\n"


def get_preprocessed_entry_point(entry_point, chief_config, worker_config, worker_count,
                                 distribution_strategy, called_from_notebook=False):
    script_lines = [
        "import os\n",
        FRAMEWORK_IMPORT,
        'os.environ["TF_KERAS_RUNNING_REMOTELY"]="1"\n',
    ]
    if distribution_strategy == "auto":
        if worker_count > 0:
            if machine_config.is_tpu_config(worker_config):
                strategy = get_tpu_cluster_resolver_fn()
                strategy.extend([
                    "resolver = wait_for_tpu_cluster_resolver_ready()\n",
                    "strategy = tfc.distribute.experimental.TPUStrategy(resolver)\n",
                ])
            else:
                strategy = ["strategy = tfc.distribute.experimental.MultiWorkerMirroredStrategy()\n"]
        elif chief_config.accelerator_count > 1:
            strategy = ["strategy = tfc.distribute.MirroredStrategy()\n"]
        else:
            strategy = ["strategy = tfc.distribute.OneDeviceStrategy(device='/gpu:0')\n"]
        script_lines.extend(strategy)
        script_lines.append("tfc.distribute.experimental_set_strategy(strategy)\n")

    if entry_point is None and not called_from_notebook:
        entry_point = sys.argv[0]

    if entry_point is not None and entry_point.endswith("py"):
        _, entry_point_file_name = os.path.split(entry_point)
        script_lines.append('exec(open("{}").read())\n'.format(entry_point_file_name))
    else:
        if called_from_notebook and entry_point is None:
            py_content = _get_live_notebook_content()
        else:
            py_content = _get_notebook_file_content(entry_point)
        for line in py_content:
            if not (line.startswith("!") or line.startswith("%") or line.startswith("#")):
                script_lines.append(line)

    _, output_file = tempfile.mkstemp(suffix=".py")
    with open(output_file, "w") as f:
        f.writelines(script_lines)
    return output_file


def _cells_to_lines(cells):
    py_content = []
    for cell in cells:
        if cell.get("cell_type") != "code":
            continue
        src = cell.get("source", [])
        if isinstance(src, str):
            src = src.splitlines(keepends=True)
        src = list(src)
        if src:
            if not src[-1].endswith("\n"):
                src[-1] += "\n"
            py_content.extend(src)
    return py_content


def _get_notebook_file_content(path):
    """Code of an ``.ipynb`` file (nbformat 4 ``cells`` or nbformat 3 ``worksheets``)."""
    try:
        with open(path) as f:
            nb = json.load(f)
    except (OSError, ValueError) as e:
        raise RuntimeError("Unable to access iPython notebook. {}".format(e))
    cells = nb.get("cells")
    if cells is None:
        cells = []
        for ws in nb.get("worksheets", []):
            for c in ws.get("cells", []):
                if c.get("cell_type") == "code":
                    cells.append({"cell_type": "code", "source": c.get("input", [])})
    return _cells_to_lines(cells)


def _get_live_notebook_content():
    """Running inside a notebook with ``entry_point=None``: Colab's ``get_ipynb`` request
    (preprocess.py:196-212) when ``google.colab`` is importable, else the notebook file named by
    ``$TFC_NOTEBOOK_PATH`` / Jupyter's ``JPY_SESSION_NAME``."""
    try:
        from google.colab import _message  # type: ignore
        response = _message.blocking_request("get_ipynb", request="", timeout_sec=200)
        if response is None:
            raise RuntimeError("Unable to get the notebook contents.")
        return _cells_to_lines(response["ipynb"]["cells"])
    except ImportError:
        pass
    path = os.environ.get("TFC_NOTEBOOK_PATH") or os.environ.get("JPY_SESSION_NAME")
    if path and os.path.exists(path):
        return _get_notebook_file_content(path)
    raise RuntimeError("Unable to get the notebook contents. Set TFC_NOTEBOOK_PATH to the .ipynb "
                       "file of the running notebook.")


def get_tpu_cluster_resolver_fn():
    """Kept for API parity (preprocess.py:215-262).  There is no TPU on this box: the generated
    function raises as soon as the wrapper runs, with a pointer to the B200 presets."""
    return [
        "def wait_for_tpu_cluster_resolver_ready():\n",
        "  raise NotImplementedError(\n",
        "      'There is no TPU on this box: use a B200 machine config '\n",
        "      '(e.g. COMMON_MACHINE_CONFIGS[\"B200_8X\"]) instead of a TPU worker_config.')\n",
    ]
