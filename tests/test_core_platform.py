"""Port of core/tests/unit/gcp_test.py (+ machine_config): table look-ups, label rules, B200 rows."""
import pytest

from cloud_b200.core import machine_config, platform as gcp

A = machine_config.AcceleratorType


def test_get_region_and_project(monkeypatch):
    assert gcp.get_region() == "us-central1"
    monkeypatch.setenv("TFC_PROJECT", "my-project")
    assert gcp.get_project_name() == "my-project"


def test_get_accelerator_type():
    assert gcp.get_accelerator_type("CPU") == "ACCELERATOR_TYPE_UNSPECIFIED"
    assert gcp.get_accelerator_type("K80") == "NVIDIA_TESLA_K80"
    assert gcp.get_accelerator_type("V100") == "NVIDIA_TESLA_V100"
    assert gcp.get_accelerator_type("TPU_V3") == "TPU_V3"
    assert gcp.get_accelerator_type("B200") == "NVIDIA_B200"


def test_get_machine_type():
    assert gcp.get_machine_type(4, 15, A.NO_ACCELERATOR) == "n1-standard-4"
    assert gcp.get_machine_type(96, 624, A.NVIDIA_TESLA_V100) == "n1-highmem-96"
    assert gcp.get_machine_type(16, 14.4, A.NVIDIA_TESLA_T4) == "n1-highcpu-16"
    assert gcp.get_machine_type(None, None, A.TPU_V3) == "cloud_tpu"


def test_table_has_the_reference_rows_and_b200():
    table = gcp._get_valid_machine_configurations()
    legacy = [t for t in table if t[2] != "B200"]
    assert len(legacy) == 199                               # gcp.py:123-406
    assert sum(1 for t in table if t[2] == "CPU") == 17
    assert (8, 30, "T4", 1) in table and (96, 624, "V100", 8) in table
    assert (64, 57.6, "K80", 8) in table and (64, 240, "K80", 8) not in table
    assert (None, None, "TPU_V2", 8) in table
    for n in (1, 2, 4, 8):
        assert (8, 30, "B200", n) in table


def test_validate_machine_configuration():
    gcp.validate_machine_configuration(4, 15, A.NVIDIA_TESLA_K80, 4)
    gcp.validate_machine_configuration(8, 30, A.NVIDIA_B200, 8)
    gcp.validate_machine_configuration(1, 1, A.TPU_V3, 8)          # cpu/memory ignored for TPU
    with pytest.raises(ValueError, match=r"Invalid machine configuration"):
        gcp.validate_machine_configuration(1, 15, A.NVIDIA_TESLA_K80, 4)
    with pytest.raises(ValueError, match=r"Invalid machine configuration"):
        gcp.validate_machine_configuration(None, None, A.TPU_V3, 4)


def test_machine_config_defaults_and_presets():
    mc = machine_config.MachineConfig()
    assert (mc.cpu_cores, mc.memory, mc.accelerator_type, mc.accelerator_count) == \
        (8, 30, A.NVIDIA_TESLA_P100, 1)
    assert len(machine_config.COMMON_MACHINE_CONFIGS) == 13 + 4
    assert machine_config.COMMON_MACHINE_CONFIGS["K80_8X"].accelerator_count == 8
    assert machine_config.is_tpu_config(machine_config.COMMON_MACHINE_CONFIGS["TPU"])
    assert not machine_config.is_tpu_config(machine_config.COMMON_MACHINE_CONFIGS["B200_8X"])
    assert not machine_config.is_tpu_config(None)
    with pytest.raises(ValueError, match=r"Invalid accelerator key provided"):
        machine_config.MachineConfig(accelerator_type="H100")
    with pytest.raises(ValueError, match=r"Invalid machine configuration"):
        machine_config.MachineConfig(cpu_cores=3)
    assert A.NVIDIA_B200 in A.all() and len(A.all()) == 9


@pytest.mark.parametrize("labels,regex", [
    ({"k%d" % i: "v" for i in range(65)}, r"too many labels"),
    ({"": "v"}, r"Label key must start with lowercase letters"),
    ({"Key": "v"}, r"Label key must start with lowercase letters"),
    ({"k": ""}, r"Label value must start with lowercase letters"),
    ({"k": "Value"}, r"Label value must start with lowercase letters"),
    ({"k" * 64: "v"}, r"Label key is too long"),
    ({"k": "v" * 64}, r"Label value is too long"),
    ({"k$": "v"}, r"Label key can only contain lowercase letters"),
    ({"k": "v$"}, r"Label value can only contain lowercase letters"),
])
def test_invalid_job_labels(labels, regex):
    with pytest.raises(ValueError, match=r"Invalid job labels"):
        gcp.validate_job_labels(labels)
    with pytest.raises(ValueError, match=regex):
        gcp.validate_job_labels(labels)


def test_valid_job_labels(capsys):
    gcp.validate_job_labels({"team": "vision", "run-1": "a_b-c"})
    gcp.validate_job_labels({})
    assert "No labels provided" in capsys.readouterr().out
