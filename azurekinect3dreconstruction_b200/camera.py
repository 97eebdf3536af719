"""o3d.camera look-alikes.

Reference call sites: GoodColorReconstruction.py:33-37, BetterVisualizer2CamWorksCOLOR.py:110-127,
Depth2PC.py:71-73 (PrimeSenseDefault).
"""
from __future__ import annotations

import enum

import numpy as np


class PinholeCameraIntrinsicParameters(enum.Enum):
    PrimeSenseDefault = 0
    Kinect2DepthCameraDefault = 1
    Kinect2ColorCameraDefault = 2


_PRESETS = {
    PinholeCameraIntrinsicParameters.PrimeSenseDefault: (640, 480, 525.0, 525.0, 319.5, 239.5),
    PinholeCameraIntrinsicParameters.Kinect2DepthCameraDefault: (512, 424, 365.456, 365.456, 254.878, 205.395),
    PinholeCameraIntrinsicParameters.Kinect2ColorCameraDefault: (1920, 1080, 1059.9718, 1059.9718, 975.7193, 545.9533),
}


class PinholeCameraIntrinsic:
    def __init__(self, *args, **kwargs):
        self.width = -1
        self.height = -1
        self.intrinsic_matrix = np.eye(3, dtype=np.float64)
        if len(args) == 1 and isinstance(args[0], PinholeCameraIntrinsicParameters):
            self.set_intrinsics(*_PRESETS[args[0]])
        elif len(args) == 1 and isinstance(args[0], PinholeCameraIntrinsic):
            o = args[0]
            self.width, self.height = o.width, o.height
            self.intrinsic_matrix = np.array(o.intrinsic_matrix, dtype=np.float64)
        elif len(args) == 3 and not kwargs:      # (width, height, intrinsic_matrix)
            self.width, self.height = int(args[0]), int(args[1])
            self.intrinsic_matrix = np.array(args[2], dtype=np.float64).reshape(3, 3)
        elif args or kwargs:
            names = ["width", "height", "fx", "fy", "cx", "cy"]
            vals = dict(zip(names, args))
            vals.update(kwargs)
            missing = [n for n in names if n not in vals]
            if missing:
                raise TypeError("PinholeCameraIntrinsic(): missing %s" % missing)
            self.set_intrinsics(*(vals[n] for n in names))

    def set_intrinsics(self, width, height, fx, fy, cx, cy):
        self.width, self.height = int(width), int(height)
        self.intrinsic_matrix = np.array([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]], dtype=np.float64)

    def get_focal_length(self):
        return float(self.intrinsic_matrix[0, 0]), float(self.intrinsic_matrix[1, 1])

    def get_principal_point(self):
        return float(self.intrinsic_matrix[0, 2]), float(self.intrinsic_matrix[1, 2])

    def get_skew(self):
        return float(self.intrinsic_matrix[0, 1])

    def is_valid(self):
        return self.width > 0 and self.height > 0

    def fxfycxcy(self):
        K = self.intrinsic_matrix
        return float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2])

    def __repr__(self):
        return "PinholeCameraIntrinsic with width = %d and height = %d.\nAccess intrinsics with intrinsic_matrix." % (
            self.width, self.height)
