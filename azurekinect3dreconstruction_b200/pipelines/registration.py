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


def _correspondence_set(corr: torch.Tensor, n: int) -> "Vector2iVector":
    """(source, target) pairs of the matched source points, ascending by source (b2r_correspondence_set)."""
    pairs = torch.empty((max(n, 1), 2), dtype=torch.int32, device=corr.device)
    cnt = C.c_int64(0)
    _lib.check(_lib.load().b2r_correspondence_set(_lib.ptr(corr.contiguous()), n, _lib.ptr(pairs), C.byref(cnt), _lib.stream_ptr()), "correspondence_set")
    return Vector2iVector(pairs[:int(cnt.value)])


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
    p2p = isinstance(estimation_method, TransformationEstimationPointToPoint)
    if p2p and estimation_method.with_scaling:
        raise RuntimeError("registration_icp: TransformationEstimationPointToPoint(with_scaling=True) is not built (the reference never scales)")
    if not p2p and not isinstance(estimation_method, TransformationEstimationPointToPlane):
        raise RuntimeError("registration_icp: TransformationEstimationPointToPoint / PointToPlane expected (coloured ICP: registration_colored_icp)")
    if not p2p and not target.has_normals():
        raise RuntimeError("[RegistrationICP] TransformationEstimationPointToPlane and TransformationEstimationColoredICP require pre-computed normal vectors for target PointCloud.")
    _lib.require_cuda()
    init = np.eye(4) if init is None else np.asarray(init, dtype=np.float64)
    N, M = len(source.points), len(target.points)
    corr = torch.empty((max(N, 1),), dtype=torch.int32, device=source.points.tensor.device)
    T = (C.c_double * 16)()
    fit, rmse = C.c_double(0), C.c_double(0)
    if p2p:
        rc = _lib.load().b2r_icp_point_to_point(_lib.ptr(source.points.tensor), N, _lib.ptr(target.points.tensor), M,
                                                float(max_correspondence_distance), _lib.dbl16(init), criteria.max_iteration,
                                                criteria.relative_fitness, criteria.relative_rmse, T, C.byref(fit), C.byref(rmse),
                                                _lib.ptr(corr), _lib.stream_ptr())
    else:
        rc = _lib.load().b2r_icp_point_to_plane(_lib.ptr(source.points.tensor), N, _lib.ptr(target.points.tensor), _lib.ptr(target.normals.tensor), M,
                                                float(max_correspondence_distance), _lib.dbl16(init), criteria.max_iteration,
                                                criteria.relative_fitness, criteria.relative_rmse, T, C.byref(fit), C.byref(rmse),
                                                _lib.ptr(corr), _lib.stream_ptr())
    _lib.check(rc, "registration_icp")
    res = RegistrationResult()
    res.transformation = np.array(T[:], dtype=np.float64).reshape(4, 4)
    res.fitness, res.inlier_rmse, res.num_iterations = float(fit.value), float(rmse.value), int(rc)
    if N and M:
        res.correspondence_set = _correspondence_set(corr, N)
    return res


class TransformationEstimationForColoredICP(TransformationEstimation):
    """Park et al. 2017 joint geometric + photometric objective (GoodColorReconstruction.py:250)."""

    def __init__(self, lambda_geometric: float = 0.968, kernel=None):
        self.lambda_geometric = float(lambda_geometric)
        if kernel is not None:
            raise RuntimeError("registration_colored_icp: robust kernels are not built (the reference passes none)")

    def __repr__(self):
        return "TransformationEstimationForColoredICP with lambda_geometric=%g" % self.lambda_geometric


def registration_colored_icp(source: PointCloud, target: PointCloud, max_correspondence_distance: float, init=None,
                             estimation_method: TransformationEstimationForColoredICP = None,
                             criteria: ICPConvergenceCriteria = None) -> RegistrationResult:
    """o3d.pipelines.registration.registration_colored_icp (GoodColorReconstruction.py:246-254).
    The target's colour gradients (hybrid search, radius 2 * max_correspondence_distance, 30 neighbours) are computed
    inside the call, then the same on-device ICP loop as point-to-plane runs with two residual rows per correspondence."""
    estimation_method = estimation_method or TransformationEstimationForColoredICP()
    criteria = criteria or ICPConvergenceCriteria()
    if not isinstance(estimation_method, TransformationEstimationForColoredICP):
        raise RuntimeError("registration_colored_icp: estimation_method must be TransformationEstimationForColoredICP")
    if not target.has_normals():
        raise RuntimeError("ColoredICP requires target pointcloud to have normals.")
    if not target.has_colors():
        raise RuntimeError("ColoredICP requires target pointcloud to have colors.")
    if not source.has_colors():
        raise RuntimeError("ColoredICP requires source pointcloud to have colors.")
    if not (0.0 <= estimation_method.lambda_geometric <= 1.0):
        raise RuntimeError("TransformationEstimationForColoredICP::lambda_geometric needs to be in range [0, 1].")
    if max_correspondence_distance <= 0.0:
        raise RuntimeError("[RegistrationICP] Invalid max_correspondence_distance.")
    _lib.require_cuda()
    init = np.eye(4) if init is None else np.asarray(init, dtype=np.float64)
    N, M = len(source.points), len(target.points)
    dev = source.points.tensor.device
    corr = torch.empty((max(N, 1),), dtype=torch.int32, device=dev)
    grad = torch.empty((max(M, 1), 3), dtype=torch.float64, device=dev)
    T = (C.c_double * 16)()
    fit, rmse = C.c_double(0), C.c_double(0)
    rc = _lib.load().b2r_icp_colored(_lib.ptr(source.points.tensor), _lib.ptr(source.colors.tensor), N, _lib.ptr(target.points.tensor),
                                     _lib.ptr(target.normals.tensor), _lib.ptr(target.colors.tensor), M, float(max_correspondence_distance),
                                     estimation_method.lambda_geometric, _lib.dbl16(init), criteria.max_iteration, criteria.relative_fitness,
                                     criteria.relative_rmse, T, C.byref(fit), C.byref(rmse), _lib.ptr(corr), _lib.ptr(grad), _lib.stream_ptr())
    _lib.check(rc, "registration_colored_icp")
    res = RegistrationResult()
    res.transformation = np.array(T[:], dtype=np.float64).reshape(4, 4)
    res.fitness, res.inlier_rmse, res.num_iterations = float(fit.value), float(rmse.value), int(rc)
    res.color_gradient = grad[:M]                    # build extension (parity hook)
    if N and M:
        res.correspondence_set = _correspondence_set(corr, N)
    return res


class Feature:
    """o3d.pipelines.registration.Feature: ``data`` is (dimension, num) like upstream; the device tensor behind it is
    point-major (num, dimension) -- ``tensor`` (build extension)."""

    def __init__(self, tensor: torch.Tensor = None):
        self.tensor = tensor if tensor is not None else torch.zeros((0, 33), dtype=torch.float64)

    @property
    def data(self) -> np.ndarray:
        return self.tensor.detach().cpu().numpy().T

    def dimension(self) -> int:
        return int(self.tensor.shape[1])

    def num(self) -> int:
        return int(self.tensor.shape[0])

    def __repr__(self):
        return "Feature class with dimension = %d and num = %d\nAccess its data via data member." % (self.dimension(), self.num())


def compute_fpfh_feature(input: PointCloud, search_param) -> Feature:
    """o3d.pipelines.registration.compute_fpfh_feature (BetterVisualizer2CamWorksCOLOR.py:211-214)."""
    from ..geometry import KDTreeSearchParamHybrid, KDTreeSearchParamKNN
    if not input.has_normals():
        raise RuntimeError("[ComputeFPFHFeature] Failed because input point cloud has no normal.")
    if isinstance(search_param, KDTreeSearchParamHybrid):
        k, radius = int(search_param.max_nn), float(search_param.radius)
    elif isinstance(search_param, KDTreeSearchParamKNN):
        k, radius = int(search_param.knn), 0.0
    else:
        raise RuntimeError("compute_fpfh_feature: KDTreeSearchParamHybrid or KDTreeSearchParamKNN expected")
    dev = _lib.require_cuda()
    n = len(input.points)
    out = torch.empty((n, 33), dtype=torch.float64, device=dev)
    rc = _lib.load().b2r_compute_fpfh(_lib.ptr(input.points.tensor), _lib.ptr(input.normals.tensor), n, k, radius, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "compute_fpfh_feature")
    return Feature(out)


class CorrespondenceChecker:
    pass


class CorrespondenceCheckerBasedOnEdgeLength(CorrespondenceChecker):
    def __init__(self, similarity_threshold: float = 0.9):
        self.similarity_threshold = float(similarity_threshold)


class CorrespondenceCheckerBasedOnDistance(CorrespondenceChecker):
    def __init__(self, distance_threshold: float):
        self.distance_threshold = float(distance_threshold)


class CorrespondenceCheckerBasedOnNormal(CorrespondenceChecker):
    def __init__(self, normal_angle_threshold: float):
        self.normal_angle_threshold = float(normal_angle_threshold)


class RANSACConvergenceCriteria:
    """RANSACConvergenceCriteria(max_iteration=100000, confidence=0.999); confidence is clamped to [0, 1] as upstream does, so
    the reference's pre-0.12 call ``RANSACConvergenceCriteria(4000000, 500)`` means "always run every iteration"."""

    def __init__(self, max_iteration: int = 100000, confidence: float = 0.999):
        self.max_iteration = int(max_iteration)
        self.confidence = max(min(float(confidence), 1.0), 0.0)


def _checker_thresholds(checkers):
    edge = dist = 0.0
    for c in checkers or []:
        if isinstance(c, CorrespondenceCheckerBasedOnEdgeLength):
            edge = c.similarity_threshold
        elif isinstance(c, CorrespondenceCheckerBasedOnDistance):
            dist = c.distance_threshold
        else:
            raise RuntimeError("registration_ransac: only the edge-length and distance checkers are built (the reference passes no other)")
    return edge, dist


def registration_ransac_based_on_correspondence(source: PointCloud, target: PointCloud, corres, max_correspondence_distance: float,
                                                estimation_method: TransformationEstimation = None, ransac_n: int = 3, checkers=None,
                                                criteria: RANSACConvergenceCriteria = None, *, seed: int = 0) -> RegistrationResult:
    """o3d.pipelines.registration.registration_ransac_based_on_correspondence.  ``seed`` (build extension) selects the draw
    sequence; the result is deterministic for a given seed."""
    estimation_method = estimation_method or TransformationEstimationPointToPoint(False)
    criteria = criteria or RANSACConvergenceCriteria()
    if not isinstance(estimation_method, TransformationEstimationPointToPoint) or getattr(estimation_method, "with_scaling", False):
        raise RuntimeError("registration_ransac: TransformationEstimationPointToPoint(False) is the estimation the reference uses and the one built")
    edge, dist = _checker_thresholds(checkers)
    dev = _lib.require_cuda()
    c = corres.tensor if hasattr(corres, "tensor") else (corres if torch.is_tensor(corres) else torch.as_tensor(np.asarray(corres)))
    c = c.to(device=dev, dtype=torch.int32).contiguous().reshape(-1, 2)
    N, M = len(source.points), len(target.points)
    T = (C.c_double * 16)()
    fit, rmse = C.c_double(0), C.c_double(0)
    rep = (C.c_int64 * 4)()
    corr = torch.full((max(N, 1),), -1, dtype=torch.int32, device=dev)
    res = RegistrationResult()
    if N and M and c.shape[0]:
        rc = _lib.load().b2r_ransac_correspondence(_lib.ptr(source.points.tensor), N, _lib.ptr(target.points.tensor), M, _lib.ptr(c), c.shape[0],
                                                   float(max_correspondence_distance), int(ransac_n), edge, dist, criteria.max_iteration,
                                                   criteria.confidence, int(seed) & 0xFFFFFFFFFFFFFFFF, T, C.byref(fit), C.byref(rmse), rep,
                                                   _lib.ptr(corr), _lib.stream_ptr())
        _lib.check(rc, "registration_ransac")
        res.transformation = np.array(T[:], dtype=np.float64).reshape(4, 4)
        res.fitness, res.inlier_rmse = float(fit.value), float(rmse.value)
        if res.fitness > 0.0:
            cc = corr[:N]
            res.correspondence_set = _correspondence_set(cc, int(cc.shape[0]))
    res.ransac_report = dict(iterations=int(rep[0]), validated=int(rep[1]), winning_iteration=int(rep[2]), inlier_correspondences=int(rep[3]))
    return res


def registration_ransac_based_on_feature_matching(source: PointCloud, target: PointCloud, source_feature: Feature, target_feature: Feature,
                                                  mutual_filter=False, max_correspondence_distance: float = None,
                                                  estimation_method: TransformationEstimation = None, ransac_n: int = 3, checkers=None,
                                                  criteria: RANSACConvergenceCriteria = None, *, seed: int = 0) -> RegistrationResult:
    """o3d.pipelines.registration.registration_ransac_based_on_feature_matching (BetterVisualizer2CamWorksCOLOR.py:230-243).
    Accepts both upstream signatures: 0.19's (``mutual_filter`` fifth) and the older one the reference was written against
    (``max_correspondence_distance`` fifth / by keyword, no mutual filter)."""
    if not isinstance(mutual_filter, (bool, np.bool_)):
        if max_correspondence_distance is None:
            max_correspondence_distance = float(mutual_filter)
        mutual_filter = False
    if max_correspondence_distance is None:
        raise TypeError("registration_ransac_based_on_feature_matching: max_correspondence_distance is required")
    if ransac_n < 3 or max_correspondence_distance <= 0.0:
        return RegistrationResult()
    dev = _lib.require_cuda()
    lib = _lib.load()
    fs, ft = source_feature.tensor.contiguous(), target_feature.tensor.contiguous()
    ns, nt = fs.shape[0], ft.shape[0]
    if ns == 0 or nt == 0:
        return RegistrationResult()
    ij = torch.empty((ns,), dtype=torch.int32, device=dev)
    _lib.check(lib.b2r_match_features(_lib.ptr(fs), ns, _lib.ptr(ft), nt, int(fs.shape[1]), _lib.ptr(ij), _lib.stream_ptr()), "feature matching")
    corres = torch.stack([torch.arange(ns, dtype=torch.int32, device=dev), ij], dim=1)
    if mutual_filter:
        ji = torch.empty((nt,), dtype=torch.int32, device=dev)
        _lib.check(lib.b2r_match_features(_lib.ptr(ft), nt, _lib.ptr(fs), ns, int(fs.shape[1]), _lib.ptr(ji), _lib.stream_ptr()), "feature matching")
        keep = ji[ij.long()] == torch.arange(ns, dtype=torch.int32, device=dev)
        mutual = corres[keep]
        if mutual.shape[0] >= ransac_n * 3:               # "Empirically mutual correspondence set should not be too small"
            corres = mutual
    return registration_ransac_based_on_correspondence(source, target, corres.contiguous(), max_correspondence_distance, estimation_method,
                                                       ransac_n, checkers, criteria, seed=seed)


# ---------------------------------------------------------------------------------------------------------------
# f3: PoseGraph + global optimisation.  Reference call sites: capture_frames.py:31-32,74-89,164-166;
# bundlefusion_reconstruction.py:179-199,232-273,292-304.  Field names and defaults follow Open3D 0.19
# (pipelines/registration/PoseGraph.h, GlobalOptimizationConvergenceCriteria.h).
# ---------------------------------------------------------------------------------------------------------------
class PoseGraphNode:
    def __init__(self, pose=None):
        self.pose = np.identity(4) if pose is None else np.array(pose, dtype=np.float64).reshape(4, 4)

    def __repr__(self):
        return "PoseGraphNode, access pose to get its current pose."


class PoseGraphEdge:
    def __init__(self, source_node_id: int = -1, target_node_id: int = -1, transformation=None, information=None,
                 uncertain: bool = False, confidence: float = 1.0):
        self.source_node_id = int(source_node_id)
        self.target_node_id = int(target_node_id)
        self.transformation = np.identity(4) if transformation is None else np.array(transformation, dtype=np.float64).reshape(4, 4)
        self.information = np.identity(6) if information is None else np.array(information, dtype=np.float64).reshape(6, 6)
        self.uncertain = bool(uncertain)
        self.confidence = float(confidence)

    def __repr__(self):
        return "PoseGraphEdge from nodes %d to %d, access transformation to get relative transformation" % (self.source_node_id, self.target_node_id)


class PoseGraph:
    def __init__(self):
        self.nodes = []
        self.edges = []

    def __repr__(self):
        return "PoseGraph with %d nodes and %d edges." % (len(self.nodes), len(self.edges))


class GlobalOptimizationOption:
    def __init__(self, max_correspondence_distance: float = 0.075, edge_prune_threshold: float = 0.25,
                 preference_loop_closure: float = 1.0, reference_node: int = -1):
        self.max_correspondence_distance = float(max_correspondence_distance)
        self.edge_prune_threshold = float(edge_prune_threshold)
        self.preference_loop_closure = float(preference_loop_closure)
        self.reference_node = int(reference_node)


class GlobalOptimizationConvergenceCriteria:
    def __init__(self):
        self.max_iteration = 100
        self.min_relative_increment = 1e-6
        self.min_relative_residual_increment = 1e-6
        self.min_right_term = 1e-6
        self.min_residual = 1e-6
        self.max_iteration_lm = 20
        self.upper_scale_factor = 2.0 / 3.0
        self.lower_scale_factor = 1.0 / 3.0


class GlobalOptimizationMethod:
    pass


class GlobalOptimizationLevenbergMarquardt(GlobalOptimizationMethod):
    pass


class GlobalOptimizationGaussNewton(GlobalOptimizationMethod):
    """Present for name compatibility; the reference only ever passes Levenberg-Marquardt."""


def global_optimization(pose_graph: PoseGraph, method: GlobalOptimizationMethod, criteria: GlobalOptimizationConvergenceCriteria,
                        option: GlobalOptimizationOption) -> None:
    """o3d.pipelines.registration.global_optimization: optimises ``pose_graph`` IN PLACE (poses updated, uncertain edges
    under the prune threshold removed, confidences refreshed).  An invalid graph is left untouched with a warning, as
    upstream does.  ``pose_graph.last_report`` (build extension) carries the iteration counts and residuals."""
    if not isinstance(method, GlobalOptimizationLevenbergMarquardt):
        raise NotImplementedError("global_optimization: only GlobalOptimizationLevenbergMarquardt is built (the reference uses no other)")
    lib = _lib.load()
    n, m = len(pose_graph.nodes), len(pose_graph.edges)
    poses = np.ascontiguousarray(np.stack([nd.pose for nd in pose_graph.nodes]).astype(np.float64)) if n else np.zeros((0, 4, 4))
    src = np.array([e.source_node_id for e in pose_graph.edges], dtype=np.int32)
    dst = np.array([e.target_node_id for e in pose_graph.edges], dtype=np.int32)
    trans = np.ascontiguousarray(np.stack([e.transformation for e in pose_graph.edges]).astype(np.float64)) if m else np.zeros((0, 4, 4))
    info = np.ascontiguousarray(np.stack([e.information for e in pose_graph.edges]).astype(np.float64)) if m else np.zeros((0, 6, 6))
    unc = np.array([1 if e.uncertain else 0 for e in pose_graph.edges], dtype=np.uint8)
    conf = np.array([e.confidence for e in pose_graph.edges], dtype=np.float64)
    keep = np.ones(m, dtype=np.uint8)
    opt = np.array([option.max_correspondence_distance, option.edge_prune_threshold, option.preference_loop_closure, option.reference_node], dtype=np.float64)
    cri = np.array([criteria.max_iteration, criteria.min_relative_increment, criteria.min_relative_residual_increment, criteria.min_right_term,
                    criteria.min_residual, criteria.max_iteration_lm, criteria.upper_scale_factor, criteria.lower_scale_factor], dtype=np.float64)
    rep = np.zeros(8, dtype=np.float64)
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a.size else None
    rc = lib.b2r_posegraph_optimize(n, p(poses), m, p(src), p(dst), p(trans), p(info), p(unc), p(conf), p(keep), p(opt), p(cri), p(rep))
    _lib.check(rc, "global_optimization")
    pose_graph.last_report = dict(valid=bool(rep[0]), iterations=(int(rep[1]), int(rep[3])), lm_steps=(int(rep[2]), int(rep[4])),
                                  residual_before=float(rep[5]), residual_after=float(rep[6]), edges_kept=int(rep[7]))
    if not rep[0]:
        import warnings
        warnings.warn("[GlobalOptimization] " + _lib.last_error())
        return
    for i, nd in enumerate(pose_graph.nodes):
        nd.pose = poses[i].copy()
    kept = []
    for k, e in enumerate(pose_graph.edges):
        if keep[k]:
            e.confidence = float(conf[k])
            kept.append(e)
    pose_graph.edges = kept
