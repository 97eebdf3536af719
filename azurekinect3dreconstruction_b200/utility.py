"""o3d.utility look-alikes: Vector3dVector / Vector3iVector / IntVector / DoubleVector.

The reference moves arrays across the Open3D boundary with these
(/root/reference/CreateMesh.py:691,720; bundlefusion_reconstruction.py:120).  Here they are thin
views over torch tensors (device-resident when a GPU is present); ``np.asarray(v)`` copies out.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


class _TensorVector:
    _dtype = torch.float64
    _cols = 3

    def __init__(self, data=None):
        if data is None:
            t = torch.empty((0, self._cols) if self._cols else (0,), dtype=self._dtype, device=_lib.default_device())
        elif isinstance(data, _TensorVector):
            t = data.tensor.clone()
        elif isinstance(data, torch.Tensor):
            t = data.to(dtype=self._dtype)
        else:
            arr = np.asarray(data)
            t = torch.as_tensor(np.ascontiguousarray(arr), dtype=self._dtype)
            t = t.to(_lib.default_device())
        if self._cols:
            if t.numel() == 0:
                t = t.reshape(0, self._cols)
            if t.dim() != 2 or t.shape[1] != self._cols:
                raise RuntimeError("expected an (N, %d) array" % self._cols)
        else:
            t = t.reshape(-1)
        self.tensor = t.contiguous()

    def __len__(self):
        return int(self.tensor.shape[0])

    def __array__(self, dtype=None, copy=None):
        a = self.tensor.detach().cpu().numpy()
        return a.astype(dtype) if dtype is not None else a

    def __getitem__(self, i):
        return self.tensor[i].detach().cpu().numpy()

    def __iter__(self):
        return iter(np.asarray(self))

    def __repr__(self):
        return "%s with %d elements." % (type(self).__name__, len(self))


class Vector3dVector(_TensorVector):
    _dtype = torch.float64
    _cols = 3


class Vector3iVector(_TensorVector):
    _dtype = torch.int32
    _cols = 3


class Vector2iVector(_TensorVector):
    _dtype = torch.int32
    _cols = 2


class IntVector(list):
    """o3d.utility.IntVector([20, 10, 5]) (bundlefusion_reconstruction.py:120)."""

    def __init__(self, data=()):
        super().__init__(int(v) for v in data)


class DoubleVector(list):
    def __init__(self, data=()):
        super().__init__(float(v) for v in data)
