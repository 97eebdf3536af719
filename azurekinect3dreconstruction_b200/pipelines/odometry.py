"""o3d.pipelines.odometry look-alike.

Reference call sites: RGBD_Odometry.py:22-35, capture_frames.py:148-151,
bundlefusion_reconstruction.py:118-129,224-227, 1camera-detectsSceneChanges.py:142,243-248.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _lib
from ..camera import PinholeCameraIntrinsic
from ..geometry import RGBDImage
from ..utility import IntVector


class OdometryOption:
    """Accepts both kwarg spellings: 0.19's depth_diff_max/depth_min/depth_max and the older
    max_depth_diff/min_depth/max_depth the reference writes (bundlefusion_reconstruction.py:122-127)."""

    def __init__(self, iteration_number_per_pyramid_level=None, depth_diff_max: float = 0.03, depth_min: float = 0.0,
                 depth_max: float = 4.0, max_depth_diff: float = None, min_depth: float = None, max_depth: float = None):
        self.iteration_number_per_pyramid_level = IntVector(iteration_number_per_pyramid_level or [20, 10, 5])
        self.depth_diff_max = float(max_depth_diff if max_depth_diff is not None else depth_diff_max)
        self.depth_min = float(min_depth if min_depth is not None else depth_min)
        self.depth_max = float(max_depth if max_depth is not None else depth_max)

    # old attribute names kept as aliases
    max_depth_diff = property(lambda s: s.depth_diff_max, lambda s, v: setattr(s, "depth_diff_max", float(v)))
    min_depth = property(lambda s: s.depth_min, lambda s, v: setattr(s, "depth_min", float(v)))
    max_depth = property(lambda s: s.depth_max, lambda s, v: setattr(s, "depth_max", float(v)))

    def __repr__(self):
        return "OdometryOption(iteration_number_per_pyramid_level=%s, depth_diff_max=%g, depth_min=%g, depth_max=%g)" % (
            list(self.iteration_number_per_pyramid_level), self.depth_diff_max, self.depth_min, self.depth_max)


class RGBDOdometryJacobian:
    kind = -1


class RGBDOdometryJacobianFromColorTerm(RGBDOdometryJacobian):
    kind = 0


class RGBDOdometryJacobianFromHybridTerm(RGBDOdometryJacobian):
    kind = 1


def compute_rgbd_odometry(rgbd_source: RGBDImage, rgbd_target: RGBDImage,
                          pinhole_camera_intrinsic: PinholeCameraIntrinsic = None, odo_init=None,
                          jacobian: RGBDOdometryJacobian = None, option: OdometryOption = None):
    """Returns (success: bool, T 4x4 f64 source->target, info 6x6 f64)."""
    intr = pinhole_camera_intrinsic or PinholeCameraIntrinsic()
    jacobian = jacobian or RGBDOdometryJacobianFromHybridTerm()
    option = option or OdometryOption()
    imgs = [rgbd_source.color.tensor, rgbd_source.depth.tensor, rgbd_target.color.tensor, rgbd_target.depth.tensor]
    shp = tuple(imgs[0].shape)
    for t in imgs:   # CheckRGBDImagePair: same size, 1 channel, 4 bytes per channel
        if t.dim() != 2 or t.dtype != torch.float32 or tuple(t.shape) != shp:
            raise RuntimeError("[ComputeRGBDOdometry] Two RGBD pairs should be same in size.")
    _lib.require_cuda()
    H, W = shp
    init = np.eye(4) if odo_init is None else np.asarray(odo_init, dtype=np.float64)
    iters = list(option.iteration_number_per_pyramid_level)
    it_arr = (C.c_int * len(iters))(*iters)
    T = (C.c_double * 16)()
    info = (C.c_double * 36)()
    ok = C.c_int(0)
    fx, fy, cx, cy = intr.fxfycxcy()
    rc = _lib.load().b2r_odometry(_lib.ptr(imgs[0].contiguous()), _lib.ptr(imgs[1].contiguous()), _lib.ptr(imgs[2].contiguous()),
                                  _lib.ptr(imgs[3].contiguous()), W, H, fx, fy, cx, cy, _lib.dbl16(init), int(jacobian.kind), it_arr,
                                  len(iters), option.depth_diff_max, option.depth_min, option.depth_max, T, info, C.byref(ok),
                                  _lib.stream_ptr())
    _lib.check(rc, "compute_rgbd_odometry")
    return bool(ok.value), np.array(T[:], dtype=np.float64).reshape(4, 4), np.array(info[:], dtype=np.float64).reshape(6, 6)


def compute_rgbd_odometry_batch(rgbd_sources, rgbd_targets, pinhole_camera_intrinsic: PinholeCameraIntrinsic = None, odo_inits=None,
                                jacobian: RGBDOdometryJacobian = None, option: OdometryOption = None, max_batch: int = 32):
    """Build extension: ``compute_rgbd_odometry`` over independent pairs (the reference loops over the consecutive frames of a
    sequence with ``odo_init = I``, RGBD_Odometry.py:26-35) through one launch sequence per ``max_batch`` pairs
    (b2r_odometry_batch).  Returns ``[(success, T, info)]``; every entry is bit-identical to the per-pair call."""
    intr = pinhole_camera_intrinsic or PinholeCameraIntrinsic()
    jacobian = jacobian or RGBDOdometryJacobianFromHybridTerm()
    option = option or OdometryOption()
    n = len(rgbd_sources)
    if len(rgbd_targets) != n:
        raise RuntimeError("[ComputeRGBDOdometry] Two RGBD pairs should be same in size.")
    if n == 0:
        return []
    shp = tuple(rgbd_sources[0].color.tensor.shape)
    imgs = [[], [], [], []]
    for s_, t_ in zip(rgbd_sources, rgbd_targets):
        four = [s_.color.tensor, s_.depth.tensor, t_.color.tensor, t_.depth.tensor]
        for k, t in enumerate(four):
            if t.dim() != 2 or t.dtype != torch.float32 or tuple(t.shape) != shp:
                raise RuntimeError("[ComputeRGBDOdometry] Two RGBD pairs should be same in size.")
            imgs[k].append(t.contiguous())
    _lib.require_cuda()
    H, W = shp
    iters = list(option.iteration_number_per_pyramid_level)
    it_arr = (C.c_int * len(iters))(*iters)
    fx, fy, cx, cy = intr.fxfycxcy()
    out = []
    for a in range(0, n, max_batch):
        b = min(n, a + max_batch)
        m = b - a
        ptrs = [(C.c_void_p * m)(*[t.data_ptr() for t in imgs[k][a:b]]) for k in range(4)]
        init = None
        if odo_inits is not None:
            init = _lib.dbl16(np.stack([np.asarray(x, dtype=np.float64) for x in odo_inits[a:b]]))
        T = (C.c_double * (16 * m))()
        info = (C.c_double * (36 * m))()
        ok = (C.c_int * m)()
        rc = _lib.load().b2r_odometry_batch(m, ptrs[0], ptrs[1], ptrs[2], ptrs[3], W, H, fx, fy, cx, cy, init, int(jacobian.kind), it_arr,
                                            len(iters), option.depth_diff_max, option.depth_min, option.depth_max, T, info, ok, _lib.stream_ptr())
        _lib.check(rc, "compute_rgbd_odometry_batch")
        Ta, Ia = np.array(T[:], dtype=np.float64).reshape(m, 4, 4), np.array(info[:], dtype=np.float64).reshape(m, 6, 6)
        out += [(bool(ok[i]), Ta[i], Ia[i]) for i in range(m)]
    return out
