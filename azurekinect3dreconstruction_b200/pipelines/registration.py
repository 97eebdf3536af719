"""o3d.pipelines.registration look-alike: point-to-plane ICP.

Reference call sites: GoodColorReconstruction.py:194-204, BetterVisualizer2CamWorksCOLOR.py:255-259,
azure-kinect-DK/Reconstruction/Reconstruction.py:129-133.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _lib
from ..geometry import PointCloud
from ..utility import Vector2iVector


class ICPConvergenceCriteria:
    def __init__(self, relative_fitness: float = 1e-6, relative_rmse: float = 1e-6, max_iteration: int = 30):
        self.relative_fitness = float(relative_fitness)
        self.relative_rmse = float(relative_rmse)
        self.max_iteration = int(max_iteration)

    def __repr__(self):
        return "ICPConvergenceCriteria(relative_fitness=%e, relative_rmse=%e, max_iteration=%d)" % (
            self.relative_fitness, self.relative_rmse, self.max_iteration)


class TransformationEstimation:
    pass


class TransformationEstimationPointToPlane(TransformationEstimation):
    def __init__(self, kernel=None):
        if kernel is not None:
            raise RuntimeError("TransformationEstimationPointToPlane: robust kernels are not supported (L2 loss only)")


class TransformationEstimationPointToPoint(TransformationEstimation):
    def __init__(self, with_scaling: bool = False):
        self.with_scaling = with_scaling


class RegistrationResult:
    def __init__(self):
        self.transformation = np.eye(4)
        self.fitness = 0.0
        self.inlier_rmse = 0.0
        self.correspondence_set = Vector2iVector()
        self.num_iterations = 0

    def __repr__(self):
        return "RegistrationResult with fitness=%e, inlier_rmse=%e, and correspondence_set size of %d\nAccess transformation to get result." % (
            self.fitness, self.inlier_rmse, len(self.correspondence_set))


def registration_icp(source: PointCloud, target: PointCloud, max_correspondence_distance: float, init=None,
                     estimation_method: TransformationEstimation = None, criteria: ICPConvergenceCriteria = None) -> RegistrationResult:
    estimation_method = estimation_method or TransformationEstimationPointToPoint()
    criteria = criteria or ICPConvergenceCriteria()
    if max_correspondence_distance <= 0.0:
        raise RuntimeError("[RegistrationICP] Invalid max_correspondence_distance.")
    if not isinstance(estimation_method, TransformationEstimationPointToPlane):
        raise RuntimeError("registration_icp: only TransformationEstimationPointToPlane is on the B200 hot path")
    if not target.has_normals():
        raise RuntimeError("[RegistrationICP] TransformationEstimationPointToPlane and TransformationEstimationColoredICP require pre-computed normal vectors for target PointCloud.")
    _lib.require_cuda()
    init = np.eye(4) if init is None else np.asarray(init, dtype=np.float64)
    N, M = len(source.points), len(target.points)
    corr = torch.empty((max(N, 1),), dtype=torch.int32, device=source.points.tensor.device)
    T = (C.c_double * 16)()
    fit, rmse = C.c_double(0), C.c_double(0)
    rc = _lib.load().b2r_icp_point_to_plane(_lib.ptr(source.points.tensor), N, _lib.ptr(target.points.tensor), _lib.ptr(target.normals.tensor), M,
                                            float(max_correspondence_distance), _lib.dbl16(init), criteria.max_iteration,
                                            criteria.relative_fitness, criteria.relative_rmse, T, C.byref(fit), C.byref(rmse),
                                            _lib.ptr(corr), _lib.stream_ptr())
    _lib.check(rc, "registration_icp")
    res = RegistrationResult()
    res.transformation = np.array(T[:], dtype=np.float64).reshape(4, 4)
    res.fitness, res.inlier_rmse, res.num_iterations = float(fit.value), float(rmse.value), int(rc)
    if N and M:
        c = corr[:N]
        src_i = torch.nonzero(c >= 0).flatten().to(torch.int32)
        res.correspondence_set = Vector2iVector(torch.stack([src_i, c[src_i.long()]], dim=1))
    return res
