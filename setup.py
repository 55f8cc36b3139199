"""Packaging for cloud_b200 (reference: src/python/setup.py:24-64 ships the version by path and the
tuner's API document as package data; same here, plus the in-tree native libraries)."""
import os

from setuptools import find_packages, setup

HERE = os.path.dirname(os.path.abspath(__file__))
version = {}
with open(os.path.join(HERE, "cloud_b200", "version.py")) as f:
    exec(f.read(), version)
deps = {}
with open(os.path.join(HERE, "dependencies.py")) as f:
    exec(f.read(), deps)

setup(
    name="cloud-b200",
    version=version["__version__"],
    description="A B200-native framework with the capabilities of TensorFlow Cloud",
    packages=find_packages(include=["cloud_b200", "cloud_b200.*"]),
    package_data={"cloud_b200.tuner": ["api/*.json"], "cloud_b200": ["_lib/*.so"]},
    python_requires=">=3.10",
    install_requires=deps["make_required_install_packages"](),
    extras_require={"tests": deps["make_required_test_packages"]()},
)
