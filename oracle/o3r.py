"""ctypes wrapper of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs --
never from the product package.  PARITY UNPINNED vs real Open3D (see oracle.cpp header).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_DIR, "liboracle.so")
_lib = None

_vp, _i, _i64, _d = C.c_void_p, C.c_int, C.c_int64, C.c_double


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    srcs = [os.path.join(_DIR, f) for f in ("oracle.cpp", "oracle_pcd.cpp")] + \
           [os.path.join(os.path.dirname(_DIR), "azurekinect3dreconstruction_b200", "csrc", "mc_tables.h")]
    if not os.path.exists(_LIB) or any(os.path.exists(f) and os.path.getmtime(f) > os.path.getmtime(_LIB) for f in srcs):
        subprocess.run(["make", "-C", _DIR], check=True, capture_output=True)
    L = C.CDLL(_LIB)
    L.o3r_depth_u16_to_f32.argtypes = [_vp, _i64, _d, _d, _vp]
    L.o3r_rgb_to_intensity.argtypes = [_vp, _i64, _vp]
    L.o3r_unproject.restype = _i64
    L.o3r_unproject.argtypes = [_vp, _i, _i, _vp, _vp, _d, _d, _d, _d, _vp, _i, _i, _vp, _vp, _vp]
    L.o3r_transform_points.argtypes = [_vp, _i64, _vp, _vp]
    L.o3r_color_gather_index.argtypes = [_vp, _i, _i, _d, _d, _d, _d, _d, _d, _i, _i, _d, _d, _d, _d, _vp, _vp]
    L.o3r_tsdf_create.restype = _vp
    L.o3r_tsdf_create.argtypes = [_d, _d, _i, _i, _i]
    L.o3r_tsdf_destroy.argtypes = [_vp]
    L.o3r_tsdf_reset.argtypes = [_vp]
    L.o3r_tsdf_set_parallel_units.argtypes = [_vp, _i]
    for n in ("o3r_tsdf_num_units", "o3r_tsdf_last_updated", "o3r_tsdf_last_touched", "o3r_tsdf_total_updated", "o3r_tsdf_total_touched",
              "o3r_tsdf_extract_points"):
        getattr(L, n).restype = _i64
        getattr(L, n).argtypes = [_vp]
    L.o3r_tsdf_integrate.argtypes = [_vp, _vp, _vp, _vp, _i, _i, _d, _d, _d, _d, _vp]
    L.o3r_tsdf_export_keys.argtypes = [_vp, _vp]
    L.o3r_tsdf_export_unit.argtypes = [_vp, _vp, _vp, _vp, _vp]
    L.o3r_tsdf_export_range.argtypes = [_vp, _i64, _i64, _vp, _vp, _vp, _vp]
    L.o3r_tsdf_import_unit.argtypes = [_vp, _vp, _vp, _vp, _vp]
    L.o3r_tsdf_extract_points_copy.argtypes = [_vp, _vp, _vp, _vp]
    L.o3r_tsdf_extract_mesh.restype = _i64
    L.o3r_tsdf_extract_mesh.argtypes = [_vp, C.POINTER(C.c_int64)]
    L.o3r_tsdf_extract_mesh_copy.argtypes = [_vp, _vp, _vp, _vp]
    L.o3r_mesh_vertex_normals.argtypes = [_vp, _i64, _vp, _i64, _vp]
    L.o3r_odometry.argtypes = [_vp, _vp, _vp, _vp, _i, _i, _d, _d, _d, _d, _vp, _i, _vp, _i, _d, _d, _d, _vp, _vp, _vp]
    L.o3r_filter.argtypes = [_vp, _i, _i, _i, _vp]
    L.o3r_downsample.argtypes = [_vp, _i, _i, _vp]
    L.o3r_correspondence.restype = _i64
    L.o3r_correspondence.argtypes = [_vp, _vp, _i, _i, _d, _d, _d, _d, _vp, _d, _vp, _i64]
    L.o3r_icp_point_to_plane.argtypes = [_vp, _i64, _vp, _vp, _i64, _d, _vp, _i, _d, _d, _vp, C.POINTER(C.c_double), C.POINTER(C.c_double), _vp]
    L.o3r_icp_point_to_point.argtypes = [_vp, _i64, _vp, _i64, _d, _vp, _i, _d, _d, _vp, C.POINTER(C.c_double), C.POINTER(C.c_double), _vp]
    L.o3r_icp_colored.argtypes = [_vp, _vp, _i64, _vp, _vp, _vp, _i64, _d, _d, _vp, _i, _d, _d, _vp, C.POINTER(C.c_double), C.POINTER(C.c_double), _vp, _vp]
    L.o3r_color_gradient.argtypes = [_vp, _vp, _vp, _i64, _d, _i, _vp]
    L.o3r_pcd_search.argtypes = [_vp, _i64, _vp, _i64, _i, _d, _vp, _vp, _vp]
    L.o3r_pcd_estimate_normals.argtypes = [_vp, _i64, _i, _d, _vp, _i, _vp]
    L.o3r_pcd_orient_normals.argtypes = [_vp, _i64, _vp, _vp]
    L.o3r_pcd_statistical_outlier.restype = _i64
    L.o3r_pcd_statistical_outlier.argtypes = [_vp, _i64, _i, _d, _vp, _vp]
    L.o3r_compute_fpfh.argtypes = [_vp, _vp, _i64, _i, _d, _vp]
    L.o3r_horn_rigid.argtypes = [_vp, _vp, _i, _vp]
    L.o3r_match_features.argtypes = [_vp, _i64, _vp, _i64, _i, _vp]
    L.o3r_ransac_correspondence.argtypes = [_vp, _i64, _vp, _i64, _vp, _i64, _d, _i, _d, _d, _i64, _d, C.c_uint64, _vp, C.POINTER(C.c_double), C.POINTER(C.c_double), _vp]
    L.o3r_pcd_voxel_down_sample.restype = _i64
    L.o3r_pcd_voxel_down_sample.argtypes = [_vp, _vp, _vp, _i64, _d, _vp, _vp, _vp, _vp]
    L.o3r_m4_inverse.argtypes = [_vp, _vp]
    L.o3r_solve6.argtypes = [_vp, _vp, _vp]
    L.o3r_vec6_to_m4.argtypes = [_vp, _vp]
    _lib = L
    return L


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def num_threads() -> int:
    return int(lib().o3r_num_threads())


def set_num_threads(n: int) -> None:
    L = lib()
    L.o3r_set_num_threads.argtypes = [C.c_int]
    L.o3r_set_num_threads.restype = None
    L.o3r_set_num_threads(int(n))


def depth_to_float(depth_u16, depth_scale=1000.0, depth_trunc=3.0):
    d = np.ascontiguousarray(depth_u16, dtype=np.uint16)
    out = np.empty(d.shape, np.float32)
    lib().o3r_depth_u16_to_f32(_p(d), d.size, depth_scale, depth_trunc, _p(out))
    return out


def rgb_to_intensity(rgb):
    c = np.ascontiguousarray(rgb, dtype=np.uint8)
    out = np.empty(c.shape[:2], np.float32)
    lib().o3r_rgb_to_intensity(_p(c), out.size, _p(out))
    return out


def unproject(depth_f32, K, extrinsic=None, rgb=None, gray=None, stride=1, valid_only=True):
    """K = (fx, fy, cx, cy). Returns (xyz (N,3) f64, col (N,3) f64 or None, src_index (N,) i32)."""
    d = np.ascontiguousarray(depth_f32, dtype=np.float32)
    H, W = d.shape
    E = _f64(np.eye(4) if extrinsic is None else extrinsic)
    cap = ((W + stride - 1) // stride) * ((H + stride - 1) // stride)
    xyz = np.empty((cap, 3), np.float64)
    has_c = rgb is not None or gray is not None
    col = np.empty((cap, 3), np.float64) if has_c else None
    idx = np.empty((cap,), np.int32)
    rgb_c = np.ascontiguousarray(rgb, dtype=np.uint8) if rgb is not None else None
    gray_c = np.ascontiguousarray(gray, dtype=np.float32) if gray is not None else None
    n = lib().o3r_unproject(_p(d), W, H, _p(rgb_c), _p(gray_c), K[0], K[1], K[2], K[3], _p(E), stride, 1 if valid_only else 0,
                            _p(xyz), _p(col), _p(idx))
    if n < 0:
        raise RuntimeError("oracle unproject failed")
    return xyz[:n], (col[:n] if has_c else None), idx[:n]


def transform_points(xyz, T, normals=None):
    p = _f64(xyz).copy()
    n = _f64(normals).copy() if normals is not None else None
    lib().o3r_transform_points(_p(p), p.shape[0], _p(_f64(T)), _p(n))
    return (p, n) if normals is not None else p


def color_gather_index(depth_u16, Kd, Wc, Hc, Kc, d2c, depth_scale=1000.0, depth_trunc=3.0):
    d = np.ascontiguousarray(depth_u16, dtype=np.uint16)
    H, W = d.shape
    out = np.empty((H, W), np.int32)
    lib().o3r_color_gather_index(_p(d), W, H, depth_scale, depth_trunc, Kd[0], Kd[1], Kd[2], Kd[3], Wc, Hc, Kc[0], Kc[1], Kc[2], Kc[3],
                                 _p(_f64(d2c)), _p(out))
    return out


class TSDF:
    def __init__(self, voxel_length, sdf_trunc, color_type=1, res=16, stride=4, parallel_units=False):
        self.h = lib().o3r_tsdf_create(voxel_length, sdf_trunc, color_type, res, stride)
        self.res, self.color_type = res, color_type
        if parallel_units:
            lib().o3r_tsdf_set_parallel_units(self.h, 1)

    def __del__(self):
        if getattr(self, "h", None):
            lib().o3r_tsdf_destroy(self.h)
            self.h = None

    def reset(self):
        lib().o3r_tsdf_reset(self.h)

    def integrate(self, depth_f32, K, extrinsic, rgb=None, gray=None):
        d = np.ascontiguousarray(depth_f32, dtype=np.float32)
        H, W = d.shape
        c = np.ascontiguousarray(rgb, dtype=np.uint8) if rgb is not None else None
        g = np.ascontiguousarray(gray, dtype=np.float32) if gray is not None else None
        rc = lib().o3r_tsdf_integrate(self.h, _p(d), _p(c), _p(g), W, H, K[0], K[1], K[2], K[3], _p(_f64(extrinsic)))
        if rc != 0:
            raise RuntimeError("[ScalableTSDFVolume::Integrate] Unsupported image format." if rc == -1 else "singular extrinsic")

    def num_units(self):
        return int(lib().o3r_tsdf_num_units(self.h))

    def last_updated(self):
        return int(lib().o3r_tsdf_last_updated(self.h))

    def last_touched(self):
        return int(lib().o3r_tsdf_last_touched(self.h))

    def total_updated(self):
        """Voxel records changed since the last reset (the roofline numerator N_upd)."""
        return int(lib().o3r_tsdf_total_updated(self.h))

    def total_touched(self):
        """(unit, frame) pairs fused since the last reset."""
        return int(lib().o3r_tsdf_total_touched(self.h))

    def export_range(self, first=0, count=None):
        """Units [first, first+count) in first-touch order: keys (n,3) i32, tsdf (n,res^3) f32, weight (n,res^3) f32,
        rgb (n,3,res^3) f64 or None."""
        n = self.num_units()
        count = n - first if count is None else min(count, n - first)
        nv = self.res ** 3
        keys = np.empty((count, 3), np.int32)
        t, w = np.empty((count, nv), np.float32), np.empty((count, nv), np.float32)
        c = np.empty((count, 3, nv), np.float64) if self.color_type else None
        if count:
            lib().o3r_tsdf_export_range(self.h, first, count, _p(keys), _p(t), _p(w), _p(c))
        return keys, t, w, c

    def keys(self):
        k = np.empty((self.num_units(), 3), np.int32)
        lib().o3r_tsdf_export_keys(self.h, _p(k))
        return k

    def unit(self, key):
        n = self.res ** 3
        t, w, c = np.empty(n, np.float32), np.empty(n, np.float32), np.empty((n, 3), np.float64)
        k = np.ascontiguousarray(key, dtype=np.int32)
        if lib().o3r_tsdf_export_unit(self.h, _p(k), _p(t), _p(w), _p(c)) != 0:
            return None
        return t, w, c

    def import_unit(self, key, tsdf, weight, rgb=None):
        k = np.ascontiguousarray(key, dtype=np.int32)
        t = np.ascontiguousarray(tsdf, dtype=np.float32)
        w = np.ascontiguousarray(weight, dtype=np.float32)
        c = _f64(rgb) if rgb is not None else None
        lib().o3r_tsdf_import_unit(self.h, _p(k), _p(t), _p(w), _p(c))

    def extract_points(self):
        n = int(lib().o3r_tsdf_extract_points(self.h))
        xyz, nrm, ids = np.empty((n, 3)), np.empty((n, 3)), np.empty((n, 4), np.int32)
        rgb = np.empty((n, 3)) if self.color_type else None
        lib().o3r_tsdf_extract_points_copy(_p(xyz), _p(rgb), _p(nrm), _p(ids))
        return xyz, rgb, nrm, ids

    def extract_mesh(self):
        nF = C.c_int64(0)
        nV = int(lib().o3r_tsdf_extract_mesh(self.h, C.byref(nF)))
        vtx, ids = np.empty((nV, 3)), np.empty((nV, 4), np.int32)
        vcol = np.empty((nV, 3)) if self.color_type else None
        tri = np.empty((int(nF.value), 3), np.int32)
        lib().o3r_tsdf_extract_mesh_copy(_p(vtx), _p(vcol), _p(tri), _p(ids))
        return vtx, vcol, tri, ids


def mesh_vertex_normals(vtx, tri):
    v, t = _f64(vtx), np.ascontiguousarray(tri, dtype=np.int32)
    out = np.empty_like(v)
    lib().o3r_mesh_vertex_normals(_p(v), v.shape[0], _p(t), t.shape[0], _p(out))
    return out


def odometry(src_gray, src_depth, tgt_gray, tgt_depth, K, init=None, jac_kind=1, iters=(20, 10, 5), depth_diff_max=0.03,
             depth_min=0.0, depth_max=4.0):
    a = [np.ascontiguousarray(x, dtype=np.float32) for x in (src_gray, src_depth, tgt_gray, tgt_depth)]
    H, W = a[0].shape
    init = _f64(np.eye(4) if init is None else init)
    it = np.ascontiguousarray(iters, dtype=np.int32)
    T, info, dbg = np.empty((4, 4)), np.empty((6, 6)), np.zeros(len(iters), np.int32)
    ok = lib().o3r_odometry(_p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), W, H, K[0], K[1], K[2], K[3], _p(init), jac_kind, _p(it), len(iters),
                            depth_diff_max, depth_min, depth_max, _p(T), _p(info), _p(dbg))
    return bool(ok), T, info, dbg


def image_filter(img, kind):
    a = np.ascontiguousarray(img, dtype=np.float32)
    out = np.empty_like(a)
    lib().o3r_filter(_p(a), a.shape[1], a.shape[0], kind, _p(out))
    return out


def image_downsample(img):
    a = np.ascontiguousarray(img, dtype=np.float32)
    out = np.empty((a.shape[0] // 2, a.shape[1] // 2), np.float32)
    lib().o3r_downsample(_p(a), a.shape[1], a.shape[0], _p(out))
    return out


def correspondence(ds, dt, K, E, depth_diff_max=0.03):
    a, b = np.ascontiguousarray(ds, dtype=np.float32), np.ascontiguousarray(dt, dtype=np.float32)
    H, W = a.shape
    out = np.empty((W * H, 4), np.int32)
    n = lib().o3r_correspondence(_p(a), _p(b), W, H, K[0], K[1], K[2], K[3], _p(_f64(E)), depth_diff_max, _p(out), W * H)
    return out[:n]


def icp_point_to_plane(src, tgt, tgt_nrm, max_dist, init=None, max_iter=30, rel_fitness=1e-6, rel_rmse=1e-6):
    s, t = _f64(src), _f64(tgt)
    n = _f64(tgt_nrm) if tgt_nrm is not None else None
    init = _f64(np.eye(4) if init is None else init)
    T = np.empty((4, 4))
    fit, rmse = C.c_double(0), C.c_double(0)
    corr = np.empty((s.shape[0],), np.int32)
    it = lib().o3r_icp_point_to_plane(_p(s), s.shape[0], _p(t), _p(n), t.shape[0], max_dist, _p(init), max_iter, rel_fitness, rel_rmse,
                                      _p(T), C.byref(fit), C.byref(rmse), _p(corr))
    if it < 0:
        raise RuntimeError("[RegistrationICP] invalid arguments (%d)" % it)
    return T, float(fit.value), float(rmse.value), corr, int(it)


def icp_point_to_point(src, tgt, max_dist, init=None, max_iter=30, rel_fitness=1e-6, rel_rmse=1e-6):
    s, t = _f64(src), _f64(tgt)
    init = _f64(np.eye(4) if init is None else init)
    T = np.empty((4, 4))
    fit, rmse = C.c_double(0), C.c_double(0)
    corr = np.empty((s.shape[0],), np.int32)
    it = lib().o3r_icp_point_to_point(_p(s), s.shape[0], _p(t), t.shape[0], max_dist, _p(init), max_iter, rel_fitness, rel_rmse,
                                      _p(T), C.byref(fit), C.byref(rmse), _p(corr))
    if it < 0:
        raise RuntimeError("[RegistrationICP] invalid arguments (%d)" % it)
    return T, float(fit.value), float(rmse.value), corr, int(it)


def icp_colored(src, src_rgb, tgt, tgt_nrm, tgt_rgb, max_dist, lambda_geometric=0.968, init=None, max_iter=30, rel_fitness=1e-6, rel_rmse=1e-6):
    """registration_colored_icp. Returns (T, fitness, inlier_rmse, corr, iterations, target colour gradients)."""
    s, t = _f64(src), _f64(tgt)
    sc = _f64(src_rgb) if src_rgb is not None else None
    tc = _f64(tgt_rgb) if tgt_rgb is not None else None
    n = _f64(tgt_nrm) if tgt_nrm is not None else None
    init = _f64(np.eye(4) if init is None else init)
    T = np.empty((4, 4))
    fit, rmse = C.c_double(0), C.c_double(0)
    corr = np.empty((s.shape[0],), np.int32)
    grad = np.empty((t.shape[0], 3))
    it = lib().o3r_icp_colored(_p(s), _p(sc), s.shape[0], _p(t), _p(n), _p(tc), t.shape[0], max_dist, lambda_geometric, _p(init), max_iter,
                               rel_fitness, rel_rmse, _p(T), C.byref(fit), C.byref(rmse), _p(corr), _p(grad))
    if it < 0:
        raise RuntimeError("[RegistrationColoredICP] invalid arguments (%d)" % it)
    return T, float(fit.value), float(rmse.value), corr, int(it), grad


def m4_inverse(m):
    out = np.empty((4, 4))
    if lib().o3r_m4_inverse(_p(_f64(m)), _p(out)) != 0:
        raise RuntimeError("singular")
    return out


def solve6(A, b):
    x = np.empty(6)
    rc = lib().o3r_solve6(_p(_f64(A)), _p(_f64(b)), _p(x))
    return x if rc == 0 else None


def vec6_to_m4(x):
    T = np.empty((4, 4))
    lib().o3r_vec6_to_m4(_p(_f64(x)), _p(T))
    return T


# ---- f1: point-cloud conditioning (oracle_pcd.cpp) ------------------------------------------------------------
def pcd_search(pts, query=None, k=20, radius=-1.0):
    """Exact brute-force neighbours: (idx [nq,k] -1 padded, d2 [nq,k], counts [nq]); query None = the points themselves."""
    p = _f64(pts)
    q = p if query is None else _f64(query)
    nq = len(q)
    idx = np.empty((nq, k), np.int32)
    d2 = np.empty((nq, k), np.float64)
    cnt = np.empty((nq,), np.int32)
    lib().o3r_pcd_search(_p(p), len(p), _p(q), nq, int(k), float(radius), _p(idx), _p(d2), _p(cnt))
    return idx, d2, cnt


def pcd_estimate_normals(pts, k=30, radius=-1.0, normals=None):
    """PointCloud::EstimateNormals (fast). Returns (normals [n,3], covariances [n,6])."""
    p = _f64(pts)
    has = normals is not None
    nrm = _f64(normals).copy() if has else np.zeros_like(p)
    cov = np.empty((len(p), 6), np.float64)
    lib().o3r_pcd_estimate_normals(_p(p), len(p), int(k), float(radius), _p(nrm), 1 if has else 0, _p(cov))
    return nrm, cov


def pcd_orient_normals(pts, normals, camera=(0.0, 0.0, 0.0)):
    p = _f64(pts)
    nrm = _f64(normals).copy()
    cam = _f64(camera)
    lib().o3r_pcd_orient_normals(_p(p), len(p), _p(cam), _p(nrm))
    return nrm


def pcd_statistical_outlier(pts, nb_neighbors=20, std_ratio=2.0):
    """PointCloud::RemoveStatisticalOutliers. Returns (keep mask [n] bool, mean neighbour distance [n])."""
    p = _f64(pts)
    keep = np.empty((len(p),), np.uint8)
    avg = np.empty((len(p),), np.float64)
    lib().o3r_pcd_statistical_outlier(_p(p), len(p), int(nb_neighbors), float(std_ratio), _p(keep), _p(avg))
    return keep.astype(bool), avg


def pcd_voxel_down_sample(pts, voxel, colors=None, normals=None):
    """PointCloud::VoxelDownSample in ascending voxel-index order. Returns (pts, colors|None, normals|None, keys [m,3])."""
    p = _f64(pts)
    c = None if colors is None else _f64(colors)
    nr = None if normals is None else _f64(normals)
    n = len(p)
    op = np.empty((n, 3)); oc = np.empty((n, 3)) if c is not None else None; on = np.empty((n, 3)) if nr is not None else None
    ok = np.empty((n, 3), np.int32)
    m = lib().o3r_pcd_voxel_down_sample(_p(p), _p(c), _p(nr), n, float(voxel), _p(op), _p(oc), _p(on), _p(ok))
    return op[:m].copy(), (oc[:m].copy() if oc is not None else None), (on[:m].copy() if on is not None else None), ok[:m].copy()


def compute_fpfh(pts, normals, max_nn=100, radius=0.0):
    """compute_fpfh_feature: (n, 33) f64, point-major (upstream's Feature.data is the transpose)."""
    p = _f64(pts)
    if normals is None:
        raise RuntimeError("[ComputeFPFHFeature] Failed because input point cloud has no normal.")
    nr = _f64(normals)
    out = np.empty((p.shape[0], 33))
    lib().o3r_compute_fpfh(_p(p), _p(nr), p.shape[0], int(max_nn), float(radius), _p(out))
    return out


def horn_rigid(s, t):
    """Rigid transform (no scale) minimising sum |R s_i + t - t_i|^2 (Horn's closed form)."""
    a, b = _f64(s), _f64(t)
    T = np.empty((4, 4))
    lib().o3r_horn_rigid(_p(a), _p(b), a.shape[0], _p(T))
    return T


def match_features(query, target):
    q, t = _f64(query), _f64(target)
    out = np.empty((q.shape[0],), np.int32)
    lib().o3r_match_features(_p(q), q.shape[0], _p(t), t.shape[0], q.shape[1], _p(out))
    return out


def ransac_correspondence(src, tgt, corres, max_dist, ransac_n=3, edge_thr=0.0, dist_thr=0.0, max_iteration=100000, confidence=0.999, seed=0):
    """Returns (T, fitness, inlier_rmse, report = (consumed, validated, winning iteration, its inlier correspondences))."""
    s, t = _f64(src), _f64(tgt)
    c = np.ascontiguousarray(corres, dtype=np.int32)
    T = np.empty((4, 4))
    fit, rmse = C.c_double(0), C.c_double(0)
    rep = np.zeros(4, np.int64)
    rc = lib().o3r_ransac_correspondence(_p(s), s.shape[0], _p(t), t.shape[0], _p(c), c.shape[0], max_dist, ransac_n, edge_thr, dist_thr,
                                         max_iteration, confidence, seed, _p(T), C.byref(fit), C.byref(rmse), _p(rep))
    if rc != 0:
        raise RuntimeError("ransac_correspondence: bad argument")
    return T, float(fit.value), float(rmse.value), tuple(int(v) for v in rep)
