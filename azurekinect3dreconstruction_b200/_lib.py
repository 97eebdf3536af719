"""ctypes binding of libb2recon.so (the C-ABI declared in include/b2recon.h).

There is NO fallback: if the library is missing it is built in-tree with nvcc; if that fails, or a
compute call is made without a CUDA device, a RuntimeError is raised (Open3D raises RuntimeError from
utility::LogError the same way, SURVEY.md 8(b) "Error conventions").
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2R_LIB") or os.path.join(_PKG, "libb2recon.so")   # B2R_LIB: a tuning variant built by _build.py

OK, ERR_BAD_ARG, ERR_CUDA, ERR_CAPACITY, ERR_FORMAT, ERR_SINGULAR, ERR_RANGE = 0, -1, -2, -3, -4, -5, -6

_vp, _i, _i64, _f, _d = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double
_pd = C.POINTER(C.c_double)
_pi64 = C.POINTER(C.c_int64)
_pi = C.POINTER(C.c_int)

# name -> (restype, argtypes); must list every function include/b2recon.h declares
SIGNATURES = {
    "b2r_version": (_i, []),
    "b2r_last_error": (C.c_char_p, []),
    "b2r_m4_inverse": (_i, [_pd, _pd]),
    "b2r_m4_mul": (None, [_pd, _pd, _pd]),
    "b2r_unit_owner": (_i, [_i, _i, _i, _i]),
    "b2r_depth_to_float": (_i, [_vp, _i64, _f, _f, _vp, _vp]),
    "b2r_depth_to_float_typed": (_i, [_vp, _i, _i64, _f, _f, _vp, _vp]),
    "b2r_rgb_to_intensity": (_i, [_vp, _i64, _i, _i, _vp, _vp]),
    "b2r_unproject_scratch_words": (_i64, [_i, _i, _i, _i, _i]),
    "b2r_unproject": (_i, [_vp, _i, _f, _f, _i, _i, _i, _vp, _i, _i, _i, _i, _vp, _pd, _pd, _pd, _vp, _i, _i, _i,
                           _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]),
    "b2r_transform_points": (_i, [_vp, _vp, _i64, _i, _pd, _vp]),
    "b2r_tsdf_create": (_vp, [_d, _d, _i, _i, _i, _i64, _i, _i]),
    "b2r_tsdf_destroy": (None, [_vp]),
    "b2r_tsdf_reset": (_i, [_vp, _vp]),
    "b2r_tsdf_voxel_length": (_d, [_vp]),
    "b2r_tsdf_sdf_trunc": (_d, [_vp]),
    "b2r_tsdf_color_type": (_i, [_vp]),
    "b2r_tsdf_capacity": (_i64, [_vp]),
    "b2r_tsdf_integrate": (_i, [_vp, _vp, _i, _f, _f, _vp, _i, _i, _vp, _i, _i, _d, _d, _d, _d, _pd, _vp]),
    "b2r_tsdf_integrate_batch": (_i, [_vp, _i, _vp, _i, _f, _f, _vp, _i, _i, _i, _i, _d, _d, _d, _d, _pd, _vp]),
    "b2r_tsdf_integrate_frames": (_i, [_vp, _i, _vp, _i, _f, _f, _vp, _i, _i, _i, _i, _d, _d, _d, _d, _pd, _vp]),
    "b2r_tsdf_integrate_host": (_i, [_vp, _i, _vp, _i, _f, _f, _vp, _i, _i, _i, _i, _d, _d, _d, _d, _pd, _vp]),
    "b2r_tsdf_stats": (_i, [_vp, _pi64, _vp]),
    "b2r_tsdf_profile": (_i, [_vp, _i]),
    "b2r_tsdf_profile_read": (_i, [_vp, _pd, _vp]),
    "b2r_tsdf_flush": (_i, [_vp, _vp]),
    "b2r_tsdf_set_batch": (_i, [_vp, _i, _vp]),
    "b2r_selftest_u8_div255": (_i, []),
    "b2r_selftest_div": (_i, [_i64, C.c_uint64, _pi64, _vp]),
    "b2r_selftest_ddiv": (_i, [_d, _i64, C.c_uint64, _pi64, _vp]),
    "b2r_tsdf_num_units": (_i64, [_vp, _vp]),
    "b2r_tsdf_export_keys": (_i, [_vp, _vp, _i64, _vp]),
    "b2r_tsdf_export_units": (_i, [_vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "b2r_tsdf_import_units": (_i, [_vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "b2r_tsdf_extract_points": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _pi64, _vp]),
    "b2r_tsdf_extract_mesh": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _pi64, _pi64, _vp]),
    "b2r_tsdf_extract_mesh_slab": (_i, [_vp, _i, _i, _i64, _vp, _vp, _vp, _vp, _i64, _i64, _pi64, _pi64, _vp]),
    "b2r_tsdf_extract_points_slab": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _i64, _pi64, _vp]),
    "b2r_peer_alloc": (_i, [_i64, C.POINTER(_vp), _vp]),
    "b2r_peer_free": (_i, [_vp]),
    "b2r_peer_open": (_i, [_vp, C.POINTER(_vp)]),
    "b2r_peer_close": (_i, [_vp]),
    "b2r_peer_copy": (_i, [_vp, _vp, _i64, _vp]),
    "b2r_stream_write32": (_i, [_vp, _vp, C.c_uint32]),
    "b2r_stream_wait32_geq": (_i, [_vp, _vp, C.c_uint32]),
    "b2r_stream_write32_multi": (_i, [_vp, _vp, _i, C.c_uint32]),
    "b2r_stream_wait32_geq_multi": (_i, [_vp, _vp, _i, C.c_uint32]),
    "b2r_peer_pull": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, C.c_uint32]),
    "b2r_tsdf_export_selected": (_i, [_vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    "b2r_tsdf_export_records": (_i, [_vp, _vp, _i64, _vp, _vp, _vp]),
    "b2r_tsdf_pool": (_vp, [_vp, _pi64]),
    "b2r_tsdf_adopt_units": (_i, [_vp, _vp, _i64, _i64, _vp]),
    "b2r_mesh_vertex_normals": (_i, [_vp, _i64, _vp, _i64, _vp, _vp]),
    "b2r_mesh_vertex_normals_noalloc": (_i, [_vp, _i64, _vp, _i64, _vp, _vp]),
    "b2r_odometry": (_i, [_vp, _vp, _vp, _vp, _i, _i, _d, _d, _d, _d, _pd, _i, _pi, _i, _d, _d, _d, _pd, _pd, _pi, _vp]),
    "b2r_odometry_batch": (_i, [_i, _vp, _vp, _vp, _vp, _i, _i, _d, _d, _d, _d, _pd, _i, _pi, _i, _d, _d, _d, _pd, _pd, _pi, _vp]),
    "b2r_image_filter": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "b2r_image_downsample": (_i, [_vp, _i, _i, _vp, _vp]),
    "b2r_knn_create": (_vp, [_vp, _i64, _d, _i, _vp]),
    "b2r_knn_destroy": (None, [_vp]),
    "b2r_knn_info": (_i, [_vp, _pd, _pi64]),
    "b2r_knn_search": (_i, [_vp, _vp, _i64, _i, _d, _vp, _vp, _vp, _vp]),
    "b2r_pcd_estimate_normals": (_i, [_vp, _vp, _i, _d, _vp, _i, _vp, _vp]),
    "b2r_pcd_orient_normals": (_i, [_vp, _i64, _pd, _vp, _vp]),
    "b2r_pcd_statistical_outlier": (_i, [_vp, _i, _d, _vp, _vp, _pi64, _vp, _vp]),
    "b2r_pcd_gather_rows": (_i, [_vp, _vp, _i64, _i, _i, _vp, _vp]),
    "b2r_pcd_voxel_down_sample": (_i, [_vp, _vp, _vp, _i64, _d, _vp, _vp, _vp, _vp, _i64, _pi64, _vp]),
    "b2r_pcd_bounds": (_i, [_vp, _i64, _pd, _vp]),
    "b2r_icp_point_to_point": (_i, [_vp, _i64, _vp, _i64, _d, _pd, _i, _d, _d, _pd, _pd, _pd, _vp, _vp]),
    "b2r_icp_colored": (_i, [_vp, _vp, _i64, _vp, _vp, _vp, _i64, _d, _d, _pd, _i, _d, _d, _pd, _pd, _pd, _vp, _vp, _vp]),
    "b2r_compute_fpfh": (_i, [_vp, _vp, _i64, _i, _d, _vp, _vp]),
    "b2r_match_features": (_i, [_vp, _i64, _vp, _i64, _i, _vp, _vp]),
    "b2r_ransac_correspondence": (_i, [_vp, _i64, _vp, _i64, _vp, _i64, _d, _i, _d, _d, _i64, _d, C.c_uint64, _pd, _pd, _pd, _pi64, _vp, _vp]),
    "b2r_knn_evaluate_transforms": (_i, [_vp, _vp, _i64, _vp, _i64, _d, _vp, _vp, _vp]),
    "b2r_posegraph_optimize": (_i, [_i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b2r_correspondence_set": (_i, [_vp, _i64, _vp, _pi64, _vp]),
    "b2r_icp_point_to_plane": (_i, [_vp, _i64, _vp, _vp, _i64, _d, _pd, _i, _d, _d, _pd, _pd, _pd, _vp, _vp]),
}

_lib = None


def load(build_if_missing: bool = True):
    """Load (building in-tree if needed) libb2recon.so and attach the C signatures."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise RuntimeError("libb2recon.so is missing (run __graft_entry__.build())")
        from . import _build
        _build.build_cuda()
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError = the C-ABI is incomplete: fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().b2r_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = "") -> int:
    if rc < 0:
        raise RuntimeError("%s%s (b2recon error %d)" % (what + ": " if what else "", last_error(), rc))
    return rc


def require_cuda() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("azurekinect3dreconstruction_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def default_device() -> torch.device:
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


def ptr(t) -> C.c_void_p:
    """Device (or host) pointer of a contiguous torch tensor, or NULL."""
    if t is None:
        return C.c_void_p(0)
    assert t.is_contiguous(), "tensor must be contiguous"
    return C.c_void_p(t.data_ptr())


def stream_ptr() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def dbl16(m) -> "C.Array":
    a = np.ascontiguousarray(np.asarray(m, dtype=np.float64).reshape(-1))
    assert a.size % 16 == 0
    return (C.c_double * a.size).from_buffer_copy(a)


def dbl4(fx, fy, cx, cy) -> "C.Array":
    return (C.c_double * 4)(float(fx), float(fy), float(cx), float(cy))


def m4_inverse(m: np.ndarray) -> np.ndarray:
    src = dbl16(m)
    dst = (C.c_double * 16)()
    rc = load().b2r_m4_inverse(src, dst)
    if rc != 0:
        raise RuntimeError("singular 4x4 matrix")
    return np.array(dst[:], dtype=np.float64).reshape(4, 4)


def m4_mul(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    dst = (C.c_double * 16)()
    load().b2r_m4_mul(dbl16(a), dbl16(b), dst)
    return np.array(dst[:], dtype=np.float64).reshape(4, 4)
