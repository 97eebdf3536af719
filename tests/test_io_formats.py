"""Host-side on-disk formats: Open3D-layout PLY / OBJ writers + readers and the reference's own
frame / trajectory / calibration files.  CPU only."""
import json

import numpy as np

import azurekinect3dreconstruction_b200 as o3d
from azurekinect3dreconstruction_b200 import io


def make_cloud(n=257, seed=0):
    rng = np.random.default_rng(seed)
    pcd = o3d.geometry.PointCloud()
    pcd.points = o3d.utility.Vector3dVector(rng.normal(size=(n, 3)))
    pcd.colors = o3d.utility.Vector3dVector(rng.uniform(0, 1, (n, 3)))
    nrm = rng.normal(size=(n, 3))
    pcd.normals = o3d.utility.Vector3dVector(nrm / np.linalg.norm(nrm, axis=1, keepdims=True))
    return pcd


def test_ply_point_cloud_round_trip_and_open3d_layout(tmp_path):
    pcd = make_cloud()
    path = str(tmp_path / "c.ply")
    assert io.write_point_cloud(path, pcd)
    raw = open(path, "rb").read()
    header = raw[:raw.index(b"end_header\n") + 11].decode()
    assert header.splitlines() == ["ply", "format binary_little_endian 1.0", "comment Created by Open3D", "element vertex 257",
                                   "property double x", "property double y", "property double z", "property double nx",
                                   "property double ny", "property double nz", "property uchar red", "property uchar green",
                                   "property uchar blue", "end_header"]
    assert len(raw) == len(header) + 257 * (6 * 8 + 3)                      # packed records, little endian
    back = io.read_point_cloud(path)
    assert np.array_equal(np.asarray(back.points), np.asarray(pcd.points))   # doubles survive bit for bit
    assert np.array_equal(np.asarray(back.normals), np.asarray(pcd.normals))
    u8 = (np.asarray(pcd.colors) * 255.0).astype(np.int64)                   # Open3D truncates colours to uchar
    assert np.array_equal(np.asarray(back.colors), u8 / 255.0)
    # ascii variant, and a cloud without attributes
    plain = o3d.geometry.PointCloud(np.asarray(pcd.points))
    p2 = str(tmp_path / "a.ply")
    assert io.write_point_cloud(p2, plain, write_ascii=True)
    b2 = io.read_point_cloud(p2)
    assert not b2.has_colors() and not b2.has_normals()
    assert np.abs(np.asarray(b2.points) - np.asarray(pcd.points)).max() < 1e-9
    assert io.write_point_cloud(str(tmp_path / "x.unknown"), pcd) is False
    assert io.read_point_cloud(str(tmp_path / "missing.ply")).is_empty()
    empty = str(tmp_path / "e.ply")
    assert io.write_point_cloud(empty, o3d.geometry.PointCloud()) and io.read_point_cloud(empty).is_empty()


def make_mesh():
    m = o3d.geometry.TriangleMesh()
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1.5]], np.float64)
    m.vertices = o3d.utility.Vector3dVector(v)
    m.triangles = o3d.utility.Vector3iVector(np.array([[0, 2, 1], [0, 1, 3], [1, 2, 3], [0, 3, 2]], np.int32))
    m.vertex_colors = o3d.utility.Vector3dVector(np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [0.5, 0.5, 0.5]], np.float64))
    n = v - v.mean(0)
    m.vertex_normals = o3d.utility.Vector3dVector(n / np.linalg.norm(n, axis=1, keepdims=True))
    return m


def test_mesh_ply_and_obj_round_trip(tmp_path):
    m = make_mesh()
    for name, kw in (("m.ply", {}), ("m_ascii.ply", {"write_ascii": True}), ("m.obj", {})):
        path = str(tmp_path / name)
        assert io.write_triangle_mesh(path, m, **kw)
        b = io.read_triangle_mesh(path)
        assert np.array_equal(np.asarray(b.triangles), np.asarray(m.triangles))
        assert np.abs(np.asarray(b.vertices) - np.asarray(m.vertices)).max() < 1e-9
        assert np.abs(np.asarray(b.vertex_normals) - np.asarray(m.vertex_normals)).max() < 1e-9
        assert np.abs(np.asarray(b.vertex_colors) - np.asarray(m.vertex_colors)).max() <= 1.0 / 255.0
    obj = open(str(tmp_path / "m.obj")).read().splitlines()
    assert obj[0].startswith("# Created by Open3D") and "f 1//1 3//3 2//2" in obj              # 1-based, v//vn
    assert sum(1 for ln in obj if ln.startswith("v ")) == 4 and sum(1 for ln in obj if ln.startswith("vn ")) == 4
    raw = open(str(tmp_path / "m.ply"), "rb").read()
    assert b"element face 4\nproperty list uchar uint vertex_indices\nend_header\n" in raw
    assert len(raw) == raw.index(b"end_header\n") + 11 + 4 * (6 * 8 + 3) + 4 * (1 + 12)
    bare = o3d.geometry.TriangleMesh()
    bare.vertices, bare.triangles = m.vertices, m.triangles
    p = str(tmp_path / "bare.obj")
    assert io.write_triangle_mesh(p, bare) and "f 1 3 2" in open(p).read()
    assert io.write_triangle_mesh(str(tmp_path / "m.stl"), m) is False


def test_reference_frame_trajectory_and_calibration_files(tmp_path):
    rng = np.random.default_rng(1)
    color = rng.integers(0, 255, (72, 128, 4), dtype=np.uint8)
    depth = rng.integers(0, 4000, (72, 128), dtype=np.uint16)
    fp = str(tmp_path / "frame_000001.npz")
    io.save_frame_npz(fp, color, depth)
    with np.load(fp) as z:                                                   # exactly the keys Depth2PC.py:190-192 writes
        assert sorted(z.files) == ["color", "depth"]
    c2, d2 = io.load_frame_npz(fp)
    assert np.array_equal(c2, color) and np.array_equal(d2, depth) and d2.dtype == np.uint16
    poses = [np.eye(4), np.array([[0.0, -1, 0, 0.1], [1, 0, 0, 0.25], [0, 0, 1, 1e-17], [0, 0, 0, 1]])]
    tp = str(tmp_path / "traj.txt")
    io.write_trajectory(tp, poses)
    lines = open(tp).read().splitlines()
    assert lines[0] == "0" and lines[5] == "1" and lines[1] == "1.0 0.0 0.0 0.0" and len(lines) == 10
    back = io.read_trajectory(tp)
    assert all(np.array_equal(a, b) for a, b in zip(back, poses))            # str(float) round-trips exactly
    Ks = [o3d.camera.PinholeCameraIntrinsic(640, 576, 504.2, 504.1, 319.8, 287.9), o3d.camera.PinholeCameraIntrinsic(1280, 720, 605.6, 605.9, 637.7, 364.3)]
    Es = [np.eye(4), poses[1]]
    cp = str(tmp_path / "calib.json")
    io.save_calibration(cp, Ks, Es, camera_serials=["000123", "000456"])
    raw = json.load(open(cp))
    assert set(raw) == {"timestamp", "num_cameras", "camera_serials", "intrinsics", "extrinsics"} and raw["num_cameras"] == 2
    assert set(raw["intrinsics"][0]) == {"width", "height", "fx", "fy", "cx", "cy"}
    K2, E2, _ = io.load_calibration(cp)
    assert np.array_equal(K2[1].intrinsic_matrix, Ks[1].intrinsic_matrix) and K2[0].width == 640
    assert np.array_equal(E2[1], Es[1])


def test_image_files_round_trip_like_the_odometry_script(tmp_path):
    """o3d.io.read_image on a colour PNG and a 16-bit depth PNG (RGBD_Odometry.py:10-13), then the RGBD pair."""
    rng = np.random.default_rng(3)
    color = rng.integers(0, 256, (48, 64, 3), dtype=np.uint8)
    depth = rng.integers(0, 4000, (48, 64), dtype=np.uint16)
    cp, dp = str(tmp_path / "c.png"), str(tmp_path / "d.png")
    assert io.write_image(cp, o3d.geometry.Image(color)) and io.write_image(dp, o3d.geometry.Image(depth))
    c2, d2 = io.read_image(cp), io.read_image(dp)
    assert np.array_equal(np.asarray(c2), color) and np.asarray(c2).dtype == np.uint8       # RGB order preserved
    assert np.array_equal(np.asarray(d2), depth) and np.asarray(d2).dtype == np.uint16
    assert (c2.width, c2.height, c2.channels, d2.bytes_per_channel) == (64, 48, 3, 2)
    try:
        io.read_image(str(tmp_path / "missing.png"))
        assert False, "expected an error"
    except RuntimeError:
        pass


def test_core_device_probe_of_the_reference():
    """GoodColorReconstruction.py:88-97: hasattr(o3d.core, 'cuda') / is_available() / Device("CUDA:0")."""
    assert hasattr(o3d, "core") and hasattr(o3d.core, "cuda")
    assert isinstance(o3d.core.cuda.is_available(), bool)
    d = o3d.core.Device("CUDA:0")
    assert d.is_cuda() and d.get_id() == 0 and repr(d) == "CUDA:0"
    assert o3d.core.Device("CPU:0") == o3d.core.Device() and o3d.core.Device("cpu:0").is_cpu()
    try:
        o3d.core.Device("TPU:0")
        assert False, "expected an error"
    except RuntimeError:
        pass
