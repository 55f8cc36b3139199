"""Port of core/tests/unit/validate_test.py: valid combinations pass; every invalid argument raises
with the reference's message prefix (matched by regex, like upstream :116-361)."""
import os

import pytest

from cloud_b200.core import machine_config, validate

CFG = machine_config.COMMON_MACHINE_CONFIGS
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(autouse=True)
def _cwd(monkeypatch):
    monkeypatch.chdir(os.path.join(HERE, "testdata"))


def _call(**over):
    kw = dict(entry_point="mnist_example_using_fit.py", requirements_txt="requirements.txt",
              distribution_strategy="auto", chief_config=CFG["K80_1X"], worker_config=CFG["K80_1X"],
              worker_count=1, region="us-central1", args=None, stream_logs=True,
              docker_image_bucket_name=None, called_from_notebook=False, job_labels={"a": "b"})
    kw.update(over)
    return validate.validate(**kw)


def test_valid_args():
    _call()
    _call(entry_point="mnist_example_using_fit.ipynb", requirements_txt=None, distribution_strategy=None,
          worker_count=0, args=["1000"], stream_logs=False, docker_image_bucket_name="bucket",
          called_from_notebook=True)
    _call(entry_point=None, requirements_txt=None, worker_count=0)        # python script as entry
    _call(entry_point=None, requirements_txt=None, docker_image_bucket_name="bucket",
          called_from_notebook=True)                                      # notebook as entry
    _call(chief_config=CFG["CPU"], worker_config=CFG["TPU"], worker_count=1, docker_base_image="custom")
    _call(chief_config=CFG["B200_8X"], worker_count=0)


@pytest.mark.parametrize("over,exc,regex", [
    (dict(entry_point="/mnist_example_using_fit.py"), ValueError, r"Invalid `entry_point`"),
    (dict(entry_point="does_not_exist.py"), ValueError, r"Invalid `entry_point`"),
    (dict(entry_point="requirements.txt"), ValueError, r"Invalid `entry_point`"),
    (dict(requirements_txt="temp.txt"), ValueError, r"Invalid `requirements_txt`"),
    (dict(distribution_strategy="MirroredStrategy"), ValueError, r"Invalid `distribution_strategy`"),
    (dict(chief_config=None), ValueError, r"Invalid `chief_config`"),
    (dict(worker_config=None), ValueError, r"Invalid `worker_config`"),
    (dict(worker_count=-1), ValueError, r"Invalid `worker_count`"),
    (dict(region=["us-region-a"]), ValueError, r"Invalid `region`"),
    (dict(args="1000"), ValueError, r"Invalid `entry_point_args`"),
    (dict(stream_logs="True"), ValueError, r"Invalid `stream_logs`"),
    (dict(called_from_notebook=True), ValueError, r"Invalid `docker_image_bucket_name`"),
    (dict(chief_config=CFG["TPU"]), ValueError, r"Invalid `chief_config`"),
    (dict(worker_config=CFG["TPU"], worker_count=2), ValueError, r"Invalid `worker_count`"),
    (dict(chief_config=CFG["CPU"], worker_config=CFG["TPU"], worker_count=1), NotImplementedError,
     r"TPUs are only supported for TF version <= 2.1.0"),
    (dict(job_labels={"A": "b"}), ValueError, r"Invalid job labels"),
])
def test_invalid_args(over, exc, regex):
    with pytest.raises(exc, match=regex):
        _call(**over)


def test_tpu_with_supported_framework_version(monkeypatch):
    from cloud_b200.utils import framework_utils
    monkeypatch.setattr(framework_utils, "get_version", lambda: "2.1")
    _call(chief_config=CFG["CPU"], worker_config=CFG["TPU"], worker_count=1)
