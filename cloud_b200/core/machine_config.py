"""Machine configurations (reference: src/python/tensorflow_cloud/core/machine_config.py:25-185).

Same enum, constructor defaults, presets and error messages as the reference, plus the hardware
this framework is written for: ``AcceleratorType.NVIDIA_B200`` and presets ``B200_{1,2,4,8}X``.
On the single 8xB200 box every accelerator kind maps to "that many local GPUs"
(see ``cloud_b200.core.launch.plan_ranks``).
"""
from __future__ import annotations

import enum

from . import platform as gcp


class AcceleratorType(enum.Enum):
    """Types of accelerators."""

    NO_ACCELERATOR = "CPU"
    NVIDIA_TESLA_K80 = "K80"
    NVIDIA_TESLA_P100 = "P100"
    NVIDIA_TESLA_V100 = "V100"
    NVIDIA_TESLA_P4 = "P4"
    NVIDIA_TESLA_T4 = "T4"
    TPU_V2 = "TPU_V2"
    TPU_V3 = "TPU_V3"
    NVIDIA_B200 = "B200"

    @classmethod
    def all(cls):
        return tuple(cls)

    @classmethod
    def validate(cls, key):
        if key not in cls.all():
            raise ValueError("Invalid accelerator key provided: %s." % key)


class MachineConfig(object):
    """Represents the configuration or type of machine to be used."""

    def __init__(self, cpu_cores=8, memory=30, accelerator_type="auto", accelerator_count=1):
        self.cpu_cores = cpu_cores
        self.memory = memory
        self.accelerator_type = accelerator_type
        self.accelerator_count = accelerator_count
        if self.accelerator_type == "auto":
            # the reference maps 'auto' to P100 (machine_config.py:82-83); the count is what matters here
            self.accelerator_type = AcceleratorType.NVIDIA_TESLA_P100
        self.validate()

    def validate(self):
        AcceleratorType.validate(self.accelerator_type)
        gcp.validate_machine_configuration(self.cpu_cores, self.memory, self.accelerator_type,
                                           self.accelerator_count)

    def __repr__(self):
        return ("MachineConfig(cpu_cores={}, memory={}, accelerator_type={}, accelerator_count={})"
                .format(self.cpu_cores, self.memory, self.accelerator_type, self.accelerator_count))


def _preset(cores, mem, accel, count):
    return MachineConfig(cpu_cores=cores, memory=mem, accelerator_type=accel, accelerator_count=count)


_A = AcceleratorType
COMMON_MACHINE_CONFIGS = {
    "CPU": _preset(4, 15, _A.NO_ACCELERATOR, 0),
    "K80_1X": _preset(8, 30, _A.NVIDIA_TESLA_K80, 1),
    "K80_4X": _preset(16, 60, _A.NVIDIA_TESLA_K80, 4),
    "K80_8X": _preset(32, 120, _A.NVIDIA_TESLA_K80, 8),
    "P100_1X": _preset(8, 30, _A.NVIDIA_TESLA_P100, 1),
    "P100_4X": _preset(16, 60, _A.NVIDIA_TESLA_P100, 4),
    "P4_1X": _preset(8, 30, _A.NVIDIA_TESLA_P4, 1),
    "P4_4X": _preset(16, 60, _A.NVIDIA_TESLA_P4, 4),
    "V100_1X": _preset(8, 30, _A.NVIDIA_TESLA_V100, 1),
    "V100_4X": _preset(16, 60, _A.NVIDIA_TESLA_V100, 4),
    "T4_1X": _preset(8, 30, _A.NVIDIA_TESLA_T4, 1),
    "T4_4X": _preset(16, 60, _A.NVIDIA_TESLA_T4, 4),
    "TPU": _preset(None, None, _A.TPU_V3, 8),
    # the hardware of this box
    "B200_1X": _preset(8, 30, _A.NVIDIA_B200, 1),
    "B200_2X": _preset(16, 60, _A.NVIDIA_B200, 2),
    "B200_4X": _preset(32, 120, _A.NVIDIA_B200, 4),
    "B200_8X": _preset(96, 360, _A.NVIDIA_B200, 8),
}


def is_tpu_config(config):
    if config:
        return config.accelerator_type in (AcceleratorType.TPU_V2, AcceleratorType.TPU_V3)
    return False
