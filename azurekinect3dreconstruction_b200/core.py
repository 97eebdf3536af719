"""The two `o3d.core` names the reference probes before choosing a device (CodeThatWorks/GoodColorReconstruction.py:88-97,
azure-kinect-DK/Reconstruction/InitialReconstruction-CUDA2.py:88-97):

    if hasattr(o3d, 'core') and hasattr(o3d.core, 'cuda') and o3d.core.cuda.is_available():
        self.device = o3d.core.Device("CUDA:0")
    else:
        self.device = o3d.core.Device("CPU:0")

Here every operation runs on the current CUDA device; `Device` is a parsed label, nothing dispatches on it (there is
no CPU path to select)."""
from __future__ import annotations

import enum

import torch


class Device:
    class DeviceType(enum.Enum):
        CPU = 0
        CUDA = 1

    def __init__(self, spec="CPU:0", device_id=None):
        if isinstance(spec, Device):
            self._type, self._id = spec._type, spec._id
            return
        if isinstance(spec, Device.DeviceType):
            self._type, self._id = spec, int(device_id or 0)
            return
        s = str(spec)
        name, _, idx = s.partition(":")
        if device_id is not None:
            idx = str(device_id)
        if name.upper() not in ("CPU", "CUDA") or (idx and not idx.isdigit()):
            raise RuntimeError(f"Invalid device string {s}. Valid device strings are like \"CPU:0\" or \"CUDA:1\"")
        self._type = Device.DeviceType.CUDA if name.upper() == "CUDA" else Device.DeviceType.CPU
        self._id = int(idx) if idx else 0

    def get_type(self):
        return self._type

    def get_id(self) -> int:
        return self._id

    def is_cuda(self) -> bool:
        return self._type == Device.DeviceType.CUDA

    def is_cpu(self) -> bool:
        return self._type == Device.DeviceType.CPU

    def __eq__(self, other):
        return isinstance(other, Device) and (self._type, self._id) == (other._type, other._id)

    def __hash__(self):
        return hash((self._type, self._id))

    def __repr__(self):
        return f"{self._type.name}:{self._id}"


class cuda:  # noqa: N801  (o3d.core.cuda is a submodule upstream; the reference only calls these two)
    @staticmethod
    def is_available() -> bool:
        return torch.cuda.is_available()

    @staticmethod
    def device_count() -> int:
        return torch.cuda.device_count()
