"""Dependency lists (reference: src/python/dependencies.py:18-38).  The reference needs
tensorflow, keras-tuner, docker and the Google client libraries; none of them are needed here."""


def make_required_install_packages():
    return ["torch>=2.4", "numpy", "cloudpickle"]


def make_required_test_packages():
    return ["pytest", "flake8", "numpy"]
