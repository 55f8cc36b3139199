"""Run a notebook without a notebook server: ``python -m cloud_b200.utils.notebook nb.ipynb``.

The reference's notebook flow - ``tfc.run()`` called from a cell with ``entry_point=None``
(core/run.py:165,249-263; core/preprocess.py:196-212; examples call_run_within_nb_on_colab.ipynb,
dogs_classification.ipynb) - needs a Jupyter or Colab kernel, and the GPU box has neither.  This
runner is the "Run all" of a kernel: it executes the code cells in order in one namespace, skips
shell (``!``) and magic (``%``) lines like the notebook converter does, and tells ``tfc.run`` which
notebook is running (``TFC_NOTEBOOK_PATH``), so a cell that calls ``tfc.run()`` submits the whole
notebook as a job, keeps going after the submission (no ``sys.exit`` inside notebooks) and the
following cells can load what the job saved.
"""
from __future__ import annotations

import argparse
import json
import os
import sys


def code_cells(path):
    """Source text of every code cell of an nbformat-4 (or -3) notebook."""
    with open(path) as f:
        nb = json.load(f)
    cells = nb.get("cells")
    if cells is None:
        cells = [{"cell_type": c.get("cell_type"), "source": c.get("input", [])}
                 for ws in nb.get("worksheets", []) for c in ws.get("cells", [])]
    out = []
    for c in cells:
        if c.get("cell_type") != "code":
            continue
        src = c.get("source", [])
        out.append(src if isinstance(src, str) else "".join(src))
    return out


def _strip_magics(source: str) -> str:
    kept = []
    for line in source.splitlines():
        stripped = line.lstrip()
        if stripped.startswith("!") or stripped.startswith("%"):
            indent = line[:len(line) - len(stripped)]
            kept.append(indent + "pass  # " + stripped)      # keeps `if x:` bodies valid
        else:
            kept.append(line)
    return "\n".join(kept) + "\n"


def run_notebook(path, namespace=None, echo=False):
    """Executes the notebook's code cells in order; returns the namespace they ran in."""
    path = os.path.abspath(path)
    ns = {"__name__": "__main__", "__file__": path} if namespace is None else namespace
    old = os.environ.get("TFC_NOTEBOOK_PATH")
    os.environ["TFC_NOTEBOOK_PATH"] = path
    try:
        for i, source in enumerate(code_cells(path)):
            if echo:
                print("In [{}]:".format(i + 1), flush=True)
            code = compile(_strip_magics(source), "{}#cell{}".format(os.path.basename(path), i + 1), "exec")
            exec(code, ns)
    finally:
        if old is None:
            os.environ.pop("TFC_NOTEBOOK_PATH", None)
        else:
            os.environ["TFC_NOTEBOOK_PATH"] = old
    return ns


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m cloud_b200.utils.notebook",
                                 description="Execute the code cells of a notebook in order.")
    ap.add_argument("notebook")
    ap.add_argument("--echo", action="store_true", help="print a cell counter before each cell")
    args = ap.parse_args(argv)
    os.chdir(os.path.dirname(os.path.abspath(args.notebook)) or ".")
    run_notebook(args.notebook, echo=args.echo)
    return 0


if __name__ == "__main__":
    sys.exit(main())
