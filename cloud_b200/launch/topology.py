"""Device-topology discovery for the launcher (north star: "local multi-process launcher with
device-topology discovery").  Answers: how many GPUs, which can reach which by P2P, is NVLS
multicast available, how much HBM, which NUMA node each GPU hangs off.

    python -m cloud_b200.launch.topology        # prints the JSON
"""
from __future__ import annotations

import dataclasses
import json
import os
import subprocess


@dataclasses.dataclass
class Topology:
    gpu_count: int = 0
    gpu_names: list = dataclasses.field(default_factory=list)
    memory_gb: list = dataclasses.field(default_factory=list)
    p2p: list = dataclasses.field(default_factory=list)          # matrix of bools
    multicast: list = dataclasses.field(default_factory=list)    # per GPU
    vmm: list = dataclasses.field(default_factory=list)
    numa: list = dataclasses.field(default_factory=list)
    cpu_count: int = 0
    all_to_all: bool = False

    def to_dict(self):
        return dataclasses.asdict(self)


def _visible_override():
    """Tests and CPU boxes can pretend to have GPUs: TFC_FAKE_GPUS=<n>."""
    v = os.environ.get("TFC_FAKE_GPUS")
    return int(v) if v else None


def gpu_count() -> int:
    """Number of GPUs this process may use, without importing torch or creating a CUDA context
    (the launcher asks this for every job): ``TFC_FAKE_GPUS``, else ``CUDA_VISIBLE_DEVICES`` if it
    is set, else NVML, else torch."""
    fake = _visible_override()
    if fake is not None:
        return fake
    visible = os.environ.get("CUDA_VISIBLE_DEVICES")
    if visible is not None:
        ids = [v for v in visible.split(",") if v.strip()]
        if not ids:
            return 0
    else:
        ids = None
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            n = int(pynvml.nvmlDeviceGetCount())
        finally:
            pynvml.nvmlShutdown()
        return min(n, len(ids)) if ids is not None else n
    except Exception:
        pass
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def discover() -> Topology:
    t = Topology(cpu_count=os.cpu_count() or 1)
    fake = _visible_override()
    if fake is not None:
        t.gpu_count = fake
        t.gpu_names = ["fake"] * fake
        t.memory_gb = [0.0] * fake
        t.p2p = [[True] * fake for _ in range(fake)]
        t.multicast = [False] * fake
        t.vmm = [False] * fake
        t.numa = [0] * fake
        t.all_to_all = fake > 1
        return t
    try:
        import torch
        if not torch.cuda.is_available():
            return t
        n = torch.cuda.device_count()
    except Exception:
        return t
    t.gpu_count = n
    from ..parallel import symm
    for i in range(n):
        props = torch.cuda.get_device_properties(i)
        t.gpu_names.append(props.name)
        t.memory_gb.append(round(props.total_memory / 2**30, 1))
        caps = symm.capabilities(i)
        t.multicast.append(caps["multicast"])
        t.vmm.append(caps["vmm"])
        t.numa.append(_numa_of(i))
    for i in range(n):
        row = []
        for j in range(n):
            if i == j:
                row.append(True)
            else:
                try:
                    row.append(bool(torch.cuda.can_device_access_peer(i, j)))
                except Exception:
                    row.append(False)
        t.p2p.append(row)
    t.all_to_all = n > 1 and all(all(r) for r in t.p2p)
    return t


def _numa_of(index: int) -> int:
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i",
                              str(index)], capture_output=True, text=True, timeout=10).stdout.strip()
        bus = out.lower()
        if bus.startswith("00000000:"):
            bus = "0000:" + bus.split(":", 1)[1]
        path = f"/sys/bus/pci/devices/{bus}/numa_node"
        if os.path.exists(path):
            with open(path) as f:
                return max(int(f.read().strip()), 0)
    except Exception:
        pass
    return 0


if __name__ == "__main__":
    print(json.dumps(discover().to_dict(), indent=1))
