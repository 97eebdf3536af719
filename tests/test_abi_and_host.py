"""CPU-side checks: the C-ABI library loads and exports every symbol include/b2recon.h declares (no compute
calls without a GPU), the host-side mirror of the Open3D API behaves like the reference expects, the
marching-cubes tables are structurally valid, and the product never reaches into oracle/."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

import azurekinect3dreconstruction_b200 as o3d
from azurekinect3dreconstruction_b200 import _lib


def header_symbols():
    src = open(os.path.join(ROOT, "include", "b2recon.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2r_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), "libb2recon.so does not export %s" % n
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert o3d.load_library().b2r_version() >= 100


def test_library_is_sm100a_and_has_no_cpu_fallback():
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True)
    if out.returncode == 0:
        assert "sm_100a" in out.stdout
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="CUDA"):
            o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, o3d.pipelines.integration.TSDFVolumeColorType.RGB8)
        with pytest.raises(RuntimeError, match="CUDA"):
            o3d.geometry.PointCloud.create_from_depth_image(o3d.geometry.Image(np.ones((4, 4), np.uint16)),
                                                            o3d.camera.PinholeCameraIntrinsic(4, 4, 1, 1, 2, 2))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "azurekinect3dreconstruction_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")) and f != "_build.py":     # _build.py only BUILDS the checker
                txt = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in txt and "o3r_" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


def test_host_pose_algebra_matches_oracle(o3r):
    rng = np.random.default_rng(0)
    for _ in range(10):
        M = np.eye(4)
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        M[:3, :3], M[:3, 3] = q, rng.normal(size=3)
        assert np.array_equal(_lib.m4_inverse(M), o3r.m4_inverse(M))      # same operation sequence on both sides
    with pytest.raises(RuntimeError):
        _lib.m4_inverse(np.zeros((4, 4)))
    A, B = rng.normal(size=(4, 4)), rng.normal(size=(4, 4))
    np.testing.assert_allclose(_lib.m4_mul(A, B), A @ B, atol=1e-14)


def test_camera_intrinsic():
    K = o3d.camera.PinholeCameraIntrinsic(o3d.camera.PinholeCameraIntrinsicParameters.PrimeSenseDefault)
    assert (K.width, K.height) == (640, 480) and K.get_focal_length() == (525.0, 525.0) and K.get_principal_point() == (319.5, 239.5)
    K2 = o3d.camera.PinholeCameraIntrinsic(1280, 720, 605.286, 605.699, 637.134, 366.758)     # GoodColorReconstruction.py:33-37
    assert K2.intrinsic_matrix[0, 2] == 637.134 and K2.intrinsic_matrix[2, 2] == 1.0
    K3 = o3d.camera.PinholeCameraIntrinsic(width=640, height=576, fx=504.2, fy=504.1, cx=319.8, cy=287.9)
    assert K3.fxfycxcy() == (504.2, 504.1, 319.8, 287.9) and K3.is_valid()
    with pytest.raises(TypeError):
        o3d.camera.PinholeCameraIntrinsic(640, 480)


def test_odometry_option_both_spellings():
    od = o3d.pipelines.odometry
    a = od.OdometryOption()
    assert list(a.iteration_number_per_pyramid_level) == [20, 10, 5] and (a.depth_diff_max, a.depth_min, a.depth_max) == (0.03, 0.0, 4.0)
    b = od.OdometryOption(iteration_number_per_pyramid_level=o3d.utility.IntVector([5, 3]), max_depth_diff=0.05, min_depth=0.1, max_depth=3.0)
    assert (b.depth_diff_max, b.depth_min, b.depth_max) == (0.05, 0.1, 3.0) and list(b.iteration_number_per_pyramid_level) == [5, 3]
    c = od.OdometryOption(depth_diff_max=0.07)
    assert c.max_depth_diff == 0.07
    assert od.RGBDOdometryJacobianFromHybridTerm().kind == 1 and od.RGBDOdometryJacobianFromColorTerm().kind == 0


def test_image_and_rgbd_host_checks():
    img = o3d.geometry.Image(np.zeros((6, 8, 3), np.uint8))
    assert (img.width, img.height, img.channels, img.bytes_per_channel) == (8, 6, 3, 1)
    d16 = o3d.geometry.Image(np.arange(48, dtype=np.uint16).reshape(6, 8))
    assert d16.bytes_per_channel == 2 and np.array_equal(np.asarray(d16), np.arange(48, dtype=np.uint16).reshape(6, 8))
    with pytest.raises(RuntimeError, match="Unsupported image format"):
        o3d.geometry.RGBDImage.create_from_color_and_depth(img, o3d.geometry.Image(np.zeros((6, 9), np.uint16)))
    # f32 depth: ConvertDepthToFloatImage still divides by depth_scale -- on the library (b2r_depth_to_float_typed);
    # without a CUDA device the shim refuses instead of computing on the host (no CPU fallback)
    f = o3d.geometry.Image(np.full((6, 8), 1500.0, np.float32))
    g = o3d.geometry.Image(np.full((6, 8), 0.5, np.float32))
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            o3d.geometry.RGBDImage.create_from_color_and_depth(g, f, 1000.0, 3.0, True)
        return
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(g, f, 1000.0, 3.0, True)
    assert np.all(np.asarray(rgbd.depth) == 1.5) and np.all(np.asarray(rgbd.color) == 0.5)
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(g, f, 1000.0, 1.5, True)
    assert np.all(np.asarray(rgbd.depth) == 0.0)                          # >= trunc -> 0
    d8 = o3d.geometry.Image(np.full((6, 8), 51, np.uint8))                # 1-channel u8 depth: / 255.0f first
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(g, d8, 0.1, 3.0, True)
    assert np.all(np.asarray(rgbd.depth) == np.float32(np.float32(51) / np.float32(255.0)) / np.float32(0.1))


def test_pointcloud_merge_semantics_and_vectors():
    import copy
    a = o3d.geometry.PointCloud(np.zeros((3, 3)))
    a.colors = o3d.utility.Vector3dVector(np.ones((3, 3)))
    b = o3d.geometry.PointCloud(np.ones((2, 3)))
    b.colors = o3d.utility.Vector3dVector(np.zeros((2, 3)))
    b.normals = o3d.utility.Vector3dVector(np.ones((2, 3)))
    c = copy.deepcopy(a)
    c += b
    assert len(c.points) == 5 and c.has_colors() and not c.has_normals()   # normals only on one side -> dropped
    assert len(a.points) == 3                                              # deepcopy is independent
    d = a + b
    assert len(d.points) == 5 and isinstance(d, o3d.geometry.PointCloud)
    e = o3d.geometry.PointCloud()
    e += b
    assert e.has_normals() and e.has_colors() and len(e.points) == 2       # empty lhs adopts the attributes
    assert np.asarray(c.points).dtype == np.float64 and len(o3d.utility.Vector3dVector()) == 0
    with pytest.raises(RuntimeError):
        o3d.utility.Vector3dVector(np.zeros((4, 2)))
    assert list(o3d.utility.IntVector([20, 10, 5])) == [20, 10, 5]
    m = o3d.geometry.TriangleMesh()
    assert m.is_empty() and not m.has_triangles()
    crit = o3d.pipelines.registration.ICPConvergenceCriteria(max_iteration=100)
    assert (crit.relative_fitness, crit.relative_rmse, crit.max_iteration) == (1e-6, 1e-6, 100)
    # paint_uniform_color (every two-camera script tints its clouds), RANSAC criteria clamping, checker classes
    assert a.paint_uniform_color([1, 0.706, 0]) is a and np.array_equal(np.asarray(a.colors), np.tile([1, 0.706, 0], (3, 1)))
    R = o3d.pipelines.registration
    assert R.RANSACConvergenceCriteria(4000000, 500).confidence == 1.0 and R.RANSACConvergenceCriteria().max_iteration == 100000
    assert R.CorrespondenceCheckerBasedOnEdgeLength(0.9).similarity_threshold == 0.9
    assert R.TransformationEstimationForColoredICP().lambda_geometric == 0.968
    assert R.GlobalOptimizationOption().reference_node == -1 and R.PoseGraphEdge().confidence == 1.0


def test_marching_cubes_tables_are_valid():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_mc_tables as g
    assert g.validate()
    # the generated header (one copy, shared by the CUDA and the oracle builds) is current
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "mc_tables.h")
        g.emit(p)
        ref = open(p).read()
    assert open(os.path.join(ROOT, "azurekinect3dreconstruction_b200/csrc/mc_tables.h")).read() == ref
    assert not os.path.exists(os.path.join(ROOT, "oracle/mc_tables.h"))


def test_synthetic_source_contract(small_kinect, tmp_path):
    from azurekinect3dreconstruction_b200 import synthetic as syn
    cap = small_kinect.render_pose(syn.orbit_trajectory(3)[0])
    di, ci = small_kinect.depth_intr, small_kinect.color_intr
    assert cap.depth.shape == (di["height"], di["width"]) and cap.depth.dtype == np.uint16      # test_device.py:23-32 shapes
    assert cap.color.shape == (ci["height"], ci["width"], 4) and cap.color.dtype == np.uint8 and np.all(cap.color[..., 3] == 255)
    assert cap.transformed_depth.shape == (ci["height"], ci["width"]) and cap.transformed_depth.dtype == np.uint16
    assert 0.9 < (cap.depth > 0).mean() <= 0.99                                               # 2 % dropout
    p = syn.save_frame_npz(str(tmp_path), 7, cap.color, cap.transformed_depth)                   # Depth2PC.py:190-192 layout
    assert p.endswith("frame_000007.npz")
    c2, d2 = syn.load_frame_npz(p)
    assert np.array_equal(c2, cap.color) and np.array_equal(d2, cap.transformed_depth)
    full = syn.SyntheticKinect()
    assert (full.depth_intr["width"], full.depth_intr["height"]) == (640, 576) and (full.color_intr["width"], full.color_intr["height"]) == (1280, 720)
    assert len(syn.orbit_trajectory(300)) == 300 and len(syn.ring_cameras(8)) == 8
