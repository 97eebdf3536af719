"""GPU parity tests: the CUDA path (through the C-ABI of libb2recon.so) vs the CPU oracle on identical
seeded inputs.  Bars (BASELINE.json north_star): bit-exact for indices / key sets / topology;
|dTSDF| <= 1e-5 (observed: bit-exact) and equal weights; vertex position <= 1e-4 * voxel;
odometry / ICP pose <= 1e-6 rad / 1e-6 m.
"""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import azurekinect3dreconstruction_b200 as o3d
from azurekinect3dreconstruction_b200 import synthetic as syn
from _helpers import image_normals

I4 = np.eye(4)


def u16(t):
    return t.view(torch.int16).cpu().numpy().view(np.uint16)


def K_of(intr):
    return (intr["fx"], intr["fy"], intr["cx"], intr["cy"])


def intrinsic_of(intr):
    return o3d.camera.PinholeCameraIntrinsic(intr["width"], intr["height"], intr["fx"], intr["fy"], intr["cx"], intr["cy"])


@pytest.fixture(scope="module")
def frames(small_kinect):
    """6 aligned (depth u16, rgb u8) frames on the small depth grid + poses (world <- camera)."""
    poses = syn.orbit_trajectory(6, yaw_sweep_deg=12.0)
    out = []
    for p in poses:
        d, c = small_kinect.render_depth_aligned_rgb(p)
        out.append((u16(d), c.cpu().numpy(), p))
    return out


# ------------------------------------------------------------------------------------------------------
# a1 / a3 / a4 / a5 / a6 unprojection
# ------------------------------------------------------------------------------------------------------
def test_rgbd_conversion_bit_exact(o3r, frames):
    d, c, _ = frames[0]
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, True)
    assert np.array_equal(np.asarray(rgbd.depth), o3r.depth_to_float(d, 1000.0, 3.0))
    assert np.array_equal(np.asarray(rgbd.color), o3r.rgb_to_intensity(c))
    with pytest.raises(RuntimeError):
        o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c[:10]), o3d.geometry.Image(d))


@pytest.mark.parametrize("extr_id", [0, 1])
def test_create_from_rgbd_image_f64_bit_exact(o3r, frames, small_kinect, extr_id):
    d, c, pose = frames[1]
    E = I4 if extr_id == 0 else np.linalg.inv(pose)
    intr = small_kinect.depth_intr
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), convert_rgb_to_intensity=False)
    pcd = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intrinsic_of(intr), E)
    xyz, col, idx = o3r.unproject(o3r.depth_to_float(d), K_of(intr), E, rgb=c)
    assert len(pcd.points) == len(xyz) > 1000
    assert np.array_equal(pcd.source_index.cpu().numpy(), idx)            # count, order, indices: bit-exact
    assert np.array_equal(np.asarray(pcd.points), xyz)                    # f64 chain: bit-exact
    assert np.array_equal(np.asarray(pcd.colors), col)
    # intensity-coloured variant (RGBD_Odometry.py:15-20)
    rgbd2 = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d))
    pcd2 = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd2, intrinsic_of(intr), E)
    _, col2, _ = o3r.unproject(o3r.depth_to_float(d), K_of(intr), E, gray=o3r.rgb_to_intensity(c))
    assert np.array_equal(np.asarray(pcd2.colors), col2)


def test_create_from_rgbd_image_f32_tolerance(o3r, frames, small_kinect):
    d, c, pose = frames[2]
    intr = small_kinect.depth_intr
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), convert_rgb_to_intensity=False)
    pcd = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intrinsic_of(intr), np.linalg.inv(pose), dtype=torch.float32)
    xyz, col, idx = o3r.unproject(o3r.depth_to_float(d), K_of(intr), np.linalg.inv(pose), rgb=c)
    assert np.array_equal(pcd.source_index.cpu().numpy(), idx)
    tol = 8 * np.spacing(np.float32(np.abs(xyz).max()))                   # 8 ulp of the largest coordinate (~2e-6 m at 3 m)
    assert np.abs(np.asarray(pcd.points) - xyz).max() <= tol
    assert np.abs(np.asarray(pcd.colors) - col).max() <= 1e-7


@pytest.mark.parametrize("stride,valid_only", [(1, True), (4, True), (3, False)])
def test_create_from_depth_image(o3r, frames, small_kinect, stride, valid_only):
    d, _, pose = frames[3]
    intr = small_kinect.depth_intr
    pcd = o3d.geometry.PointCloud.create_from_depth_image(o3d.geometry.Image(d), intrinsic_of(intr), np.linalg.inv(pose), depth_scale=1000.0,
                                                          depth_trunc=2.5, stride=stride, project_valid_depth_only=valid_only)
    xyz, _, idx = o3r.unproject(o3r.depth_to_float(d, 1000.0, 2.5), K_of(intr), np.linalg.inv(pose), stride=stride, valid_only=valid_only)
    assert np.array_equal(pcd.source_index.cpu().numpy(), idx)
    assert np.array_equal(np.asarray(pcd.points), xyz, equal_nan=True)


def test_unproject_edge_cases(o3r, small_kinect):
    intr = small_kinect.depth_intr
    W, H = intr["width"], intr["height"]
    empty = o3d.geometry.PointCloud.create_from_depth_image(o3d.geometry.Image(np.zeros((H, W), np.uint16)), intrinsic_of(intr))
    assert len(empty.points) == 0 and empty.is_empty()
    # ragged size (not a multiple of 8 / of the chunk) and a single valid pixel in the last position
    d = np.zeros((37, 53), np.uint16)
    d[-1, -1] = 1500
    d[0, 0] = 700
    K = o3d.camera.PinholeCameraIntrinsic(53, 37, 60.0, 61.0, 26.0, 18.0)
    pcd = o3d.geometry.PointCloud.create_from_depth_image(o3d.geometry.Image(d), K)
    xyz, _, idx = o3r.unproject(o3r.depth_to_float(d, 1000.0, 1000.0), (60.0, 61.0, 26.0, 18.0))
    assert np.array_equal(pcd.source_index.cpu().numpy(), idx) and np.array_equal(np.asarray(pcd.points), xyz)
    # maximum depth value, f32 input used as-is
    f = np.full((37, 53), 65.535, np.float32)
    pcd = o3d.geometry.PointCloud.create_from_depth_image(o3d.geometry.Image(f), K)
    xyz, _, _ = o3r.unproject(f, (60.0, 61.0, 26.0, 18.0))
    assert np.array_equal(np.asarray(pcd.points), xyz)


def test_kinect_capture_gather_and_two_camera_fusion(o3r, small_kinect):
    """Config 1/2: native-grid u16 depth + BGRA colour in its own camera; gather indices bit-exact;
    the flip and the per-camera extrinsic are folded into the same launch (BetterVisualizer2CamWorksCOLOR.py:534,634-640)."""
    cap = small_kinect.render_pose(syn.orbit_trajectory(3)[1])
    di, ci, d2c = small_kinect.depth_intr, small_kinect.color_intr, small_kinect.d2c
    flip = np.diag([1.0, -1.0, -1.0, 1.0])
    cam1 = syn.make_pose(syn.rot_y(math.radians(35.0)), (0.8, 0.0, 0.3))
    gi_ref = o3r.color_gather_index(cap.depth, K_of(di), ci["width"], ci["height"], K_of(ci), d2c)
    merged = o3d.geometry.PointCloud()
    for ext in (I4, cam1):
        post = ext @ flip
        pcd = o3d.geometry.PointCloud.create_from_kinect_capture(cap.depth, cap.color, intrinsic_of(di), intrinsic_of(ci), d2c,
                                                                 post_transform=post, dtype=torch.float64)
        src = pcd.source_index.cpu().numpy()
        xyz, _, idx = o3r.unproject(o3r.depth_to_float(cap.depth), K_of(di))
        assert np.array_equal(src, idx)
        assert np.array_equal(pcd.gather_index.cpu().numpy(), gi_ref.reshape(-1)[idx])     # gather index: bit-exact
        ref = o3r.transform_points(xyz, post)
        assert np.abs(np.asarray(pcd.points) - ref).max() < 1e-12
        g = gi_ref.reshape(-1)[idx]
        bgra = cap.color.reshape(-1, 4)
        rgb_ref = np.where((g >= 0)[:, None], bgra[np.maximum(g, 0)][:, [2, 1, 0]] / 255.0, 0.0)
        assert np.array_equal(np.asarray(pcd.colors), rgb_ref)
        merged += pcd
    assert len(merged.points) == 2 * len(idx) and merged.has_colors()
    assert (gi_ref >= 0).mean() > 0.5


def test_unproject_f32_single_pass_fast_path(o3r):
    """The f32 throughput path on a FAST-eligible grid (W % 32 == 0, W >= 256): one kernel, lane-interleaved pixel
    mapping, raster offsets by decoupled look-back.  Counts, raster order, source and gather indices bit-exact vs the
    oracle; colours exactly RN(c/255); coordinates within 8 ulp of the largest coordinate of the f64 chain.  A batch large
    enough for 8-chunk runs (and a ragged batch of different frames) exercises the look-back across runs."""
    from azurekinect3dreconstruction_b200.geometry import unproject
    d = dict(width=320, height=288, fx=504.2 / 2, fy=504.1 / 2, cx=319.8 / 2, cy=287.9 / 2)
    c = dict(width=640, height=360, fx=605.6 / 2, fy=605.9 / 2, cx=637.7 / 2, cy=364.3 / 2)
    kin = syn.SyntheticKinect(depth_intrinsics=d, color_intrinsics=c, seed=3, dropout=0.05)
    poses = syn.orbit_trajectory(3, yaw_sweep_deg=20.0)
    caps = [kin.render_pose(p, want_transformed_depth=False) for p in poses]
    flip = np.diag([1.0, -1.0, -1.0, 1.0])
    Kd, Kc = intrinsic_of(d), intrinsic_of(c)

    def run(idx_list, want_index):
        depth = torch.stack([torch.from_numpy(np.ascontiguousarray(caps[i].depth).view(np.int16)) for i in idx_list]).cuda().view(torch.uint16)
        color = torch.stack([torch.as_tensor(caps[i].color) for i in idx_list]).cuda()
        T = np.stack([flip @ poses[i] for i in idx_list])
        return unproject(depth, Kd, depth_is_u16=True, depth_scale=1000.0, depth_trunc=3.0, color=color, color_bgr=True, color_intrinsic=Kc,
                         depth_to_color=kin.d2c, transforms=T, dtype=torch.float32, want_index=want_index, want_gather_index=want_index)

    ref = []
    for i in range(3):
        xyz, _, idx = o3r.unproject(o3r.depth_to_float(caps[i].depth), K_of(d))
        gi = o3r.color_gather_index(caps[i].depth, K_of(d), c["width"], c["height"], K_of(c), kin.d2c).reshape(-1)[idx]
        bgra = caps[i].color.reshape(-1, 4)
        rgb = np.where((gi >= 0)[:, None], (bgra[np.maximum(gi, 0)][:, [2, 1, 0]].astype(np.float32) / np.float32(255.0)), np.float32(0))
        ref.append((o3r.transform_points(xyz, flip @ poses[i]), idx, gi, rgb))
        assert 1000 < len(idx) < d["width"] * d["height"] and (gi >= 0).mean() > 0.5

    def check(r, slot, i, with_index):
        xyz, idx, gi, rgb = ref[i]
        n = int(r["counts"][slot])
        assert n == len(idx)
        if with_index:
            assert np.array_equal(r["src_index"][slot, :n].cpu().numpy(), idx)
            assert np.array_equal(r["gather_index"][slot, :n].cpu().numpy(), gi)
        assert np.array_equal(r["rgb"][slot, :n].cpu().numpy(), rgb)
        tol = 8 * np.spacing(np.float32(np.abs(xyz).max()))
        assert np.abs(r["xyz"][slot, :n].cpu().numpy().astype(np.float64) - xyz).max() <= tol

    r = run([0, 1, 2], True)                    # run length 1: look-back over 90 single-chunk runs per frame
    for i in range(3):
        check(r, i, i, True)
    r = run([0, 1, 2], False)                   # no index staging (the throughput configuration)
    for i in range(3):
        check(r, i, i, False)
    big = [i % 3 for i in range(190)]           # 190 x 90 chunks >= 8 x 2048: runs of 8 chunks (the last run of a frame has 2)
    r = run(big, False)
    for slot in (0, 1, 2, 94, 188, 189):
        check(r, slot, big[slot], False)
    cnt = r["counts"].cpu().numpy()
    assert all(cnt[s] == len(ref[big[s]][1]) for s in range(190))


def test_transform_and_batch(o3r, frames, small_kinect):
    intr = small_kinect.depth_intr
    B = 4
    depth = torch.from_numpy(np.stack([frames[i][0] for i in range(B)]).view(np.int16)).view(torch.uint16).cuda()
    color = torch.stack([torch.from_numpy(frames[i][1]) for i in range(B)]).cuda()
    T = np.stack([frames[i][2] for i in range(B)])
    from azurekinect3dreconstruction_b200.geometry import unproject
    r = unproject(depth, intrinsic_of(intr), depth_is_u16=True, depth_scale=1000.0, depth_trunc=3.0, color=color, transforms=T,
                  dtype=torch.float64, want_index=True)
    for i in range(B):
        xyz, col, idx = o3r.unproject(o3r.depth_to_float(frames[i][0]), K_of(intr), np.linalg.inv(T[i]), rgb=frames[i][1])
        n = int(r["counts"][i])
        assert n == len(idx) and np.array_equal(r["src_index"][i, :n].cpu().numpy(), idx)
        assert np.abs(r["xyz"][i, :n].cpu().numpy() - xyz).max() < 1e-12      # T vs inv(inv(T)) rounding only
        assert np.array_equal(r["rgb"][i, :n].cpu().numpy(), col)
    pcd = o3d.geometry.PointCloud(r["xyz"][0, :int(r["counts"][0])].clone())
    M = syn.make_pose(syn.rot_x(0.2), (1, 2, 3))
    ref = o3r.transform_points(np.asarray(pcd.points), M)
    assert pcd.transform(M) is pcd and np.array_equal(np.asarray(pcd.points), ref)


# ------------------------------------------------------------------------------------------------------
# a7 / a8 / a9 TSDF
# ------------------------------------------------------------------------------------------------------
def run_both(o3r, frames, intr, vl, trunc, color_type, n, capacity=6000, batch=None):
    vol = o3d.pipelines.integration.ScalableTSDFVolume(vl, trunc, o3d.pipelines.integration.TSDFVolumeColorType(color_type), capacity_units=capacity)
    if batch:
        vol.set_batch(batch)       # frames fused per pass over the touched units (results must not depend on it)
    ref = o3r.TSDF(vl, trunc, color_type, parallel_units=True)
    K = intrinsic_of(intr)
    for d, c, pose in frames[:n]:
        E = np.linalg.inv(pose)
        if color_type == 2:
            rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, True)
            ref.integrate(o3r.depth_to_float(d), K_of(intr), E, gray=o3r.rgb_to_intensity(c))
        else:
            rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, False)
            ref.integrate(o3r.depth_to_float(d), K_of(intr), E, rgb=c if color_type == 1 else None)
        vol.integrate(rgbd, K, E)
    return vol, ref


def compare_volumes(vol, ref, color_type):
    keys, tsdf, wgt, rgb = vol.export_units()
    keys = keys.cpu().numpy()
    rk = ref.keys()
    assert {tuple(k) for k in keys.tolist()} == {tuple(k) for k in rk.tolist()}      # block allocation: bit-exact set
    assert len(keys) == len(rk)
    tsdf, wgt = tsdf.cpu().numpy(), wgt.cpu().numpy()
    rgb = rgb.cpu().numpy() if rgb is not None else None
    max_dt, n_upd, exact = 0.0, 0, True
    for i, k in enumerate(keys):
        t, w, c = ref.unit(k)
        assert np.array_equal(wgt[i], w), "weights differ in unit %s" % (k,)
        max_dt = max(max_dt, float(np.abs(tsdf[i] - t).max()))
        exact = exact and np.array_equal(tsdf[i], t)
        n_upd += int((w > 0).sum())
        if color_type:
            assert np.abs(rgb[i].T - c).max() < 1e-3                                      # f32 vs upstream's f64 running mean, 0..255 scale
    assert max_dt <= 1e-5
    return max_dt, exact, n_upd


@pytest.mark.parametrize("color_type,batch", [(0, 1), (1, 1), (2, 1), (1, 3), (1, None), (0, 16), (2, 2)])
def test_tsdf_integrate_parity(o3r, frames, small_kinect, color_type, batch):
    vol, ref = run_both(o3r, frames, small_kinect.depth_intr, 0.01, 0.04, color_type, 4, batch=batch)
    max_dt, exact, n_upd = compare_volumes(vol, ref, color_type)
    assert n_upd > 100000
    assert exact, "TSDF expected bit-exact (max diff %g)" % max_dt
    st = vol.stats()
    assert st["num_units"] == ref.num_units() and st["errors"] == 0
    if batch == 1:
        assert st["last_touched"] == ref.last_touched()


def test_shared_reciprocal_division_is_ieee_exact():
    """The fusion kernel's divisions share reciprocals (Markstein correction); they must equal div.rn bit for bit."""
    import ctypes as C
    from azurekinect3dreconstruction_b200 import _lib
    bad = C.c_int64(-1)
    _lib.check(_lib.load().b2r_selftest_div(1 << 28, 12345, C.byref(bad), _lib.stream_ptr()))
    assert bad.value == 0


def test_allocation_f64_division_by_call_constants_is_ieee_exact():
    """a8's samples divide by fx, fy and the unit length through one host-rounded reciprocal each (K1); == div.rn.f64."""
    import ctypes as C
    from azurekinect3dreconstruction_b200 import _lib
    for b in (504.2, 504.1, 605.6, 0.064, 0.032, 0.64, 4.0 / 512.0 * 16, 525.0, 1.0, 3.0, 1.0 / 3.0):
        bad = C.c_int64(-1)
        _lib.check(_lib.load().b2r_selftest_ddiv(b, 1 << 24, 777, C.byref(bad), _lib.stream_ptr()))
        assert bad.value == 0, b


def test_tsdf_batch_size_invariance(frames, small_kinect):
    """Deferred multi-frame fusion == frame-by-frame fusion, bit for bit, on every plane (incl. colour)."""
    intr = small_kinect.depth_intr
    K = intrinsic_of(intr)
    n = 6
    depth = torch.from_numpy(np.stack([frames[i][0] for i in range(n)]).view(np.int16)).view(torch.uint16).cuda()
    color = torch.stack([torch.from_numpy(frames[i][1]) for i in range(n)]).cuda()
    E = np.stack([np.linalg.inv(frames[i][2]) for i in range(n)])
    out = []
    for batch in (1, 4, 16, 32):
        v = o3d.pipelines.integration.ScalableTSDFVolume(0.004, 0.02, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, capacity_units=30000)
        v.set_batch(batch)
        v.integrate_raw(depth[:2], color[:2], K, E[:2])
        v.integrate_raw(depth[2:], color[2:], K, E[2:])
        keys, tsdf, wgt, rgb = v.export_units()
        order = np.lexsort(keys.cpu().numpy().T[::-1])
        out.append((keys.cpu().numpy()[order], tsdf.cpu().numpy()[order], wgt.cpu().numpy()[order], rgb.cpu().numpy()[order], v.stats()))
    for o in out[1:]:
        for a, b in zip(out[0][:4], o[:4]):
            assert np.array_equal(a, b)
        assert o[4]["updated_voxels"] == out[0][4]["updated_voxels"] and o[4]["touched_units"] == out[0][4]["touched_units"]


def test_tsdf_fine_voxels_and_reset(o3r, frames, small_kinect):
    vol, ref = run_both(o3r, frames, small_kinect.depth_intr, 0.004, 0.02, 1, 2, capacity=20000)
    _, exact, _ = compare_volumes(vol, ref, 1)
    assert exact
    upd_gpu = vol.stats()["updated_voxels"]
    assert upd_gpu > 0
    vol.reset()
    assert vol.num_units() == 0 and vol.stats()["updated_voxels"] == 0
    # re-integration after reset reproduces the same volume (bundlefusion_reconstruction.py:309-315)
    K = intrinsic_of(small_kinect.depth_intr)
    for d, c, pose in frames[:2]:
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, False)
        vol.integrate(rgbd, K, np.linalg.inv(pose))
    _, exact, _ = compare_volumes(vol, ref, 1)
    assert exact and vol.stats()["updated_voxels"] == upd_gpu


def test_tsdf_raw_batch_and_host_paths_match(o3r, frames, small_kinect):
    intr = small_kinect.depth_intr
    K = intrinsic_of(intr)
    n = 5
    depth = torch.from_numpy(np.stack([frames[i][0] for i in range(n)]).view(np.int16)).view(torch.uint16)
    color = torch.stack([torch.from_numpy(frames[i][1]) for i in range(n)])
    E = np.stack([np.linalg.inv(frames[i][2]) for i in range(n)])
    ref = o3r.TSDF(0.01, 0.04, 1, parallel_units=True)
    for i in range(n):
        ref.integrate(o3r.depth_to_float(frames[i][0]), K_of(intr), E[i], rgb=frames[i][1])
    mk = lambda: o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, capacity_units=6000)
    v_dev = mk()
    v_dev.integrate_raw(depth.cuda(), color.cuda(), K, E)
    assert compare_volumes(v_dev, ref, 1)[1]
    v_host = mk()
    v_host.integrate_raw(depth.pin_memory(), color.pin_memory(), K, E)       # H2D ring overlapped with fusion
    torch.cuda.synchronize()
    assert compare_volumes(v_host, ref, 1)[1]


def test_tsdf_errors(frames, small_kinect):
    d, c, pose = frames[0]
    intr = small_kinect.depth_intr
    vol = o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, capacity_units=64)
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), convert_rgb_to_intensity=False)
    with pytest.raises(RuntimeError, match="Unsupported image format"):
        vol.integrate(rgbd, o3d.camera.PinholeCameraIntrinsic(o3d.camera.PinholeCameraIntrinsicParameters.PrimeSenseDefault), I4)
    gray = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d))
    with pytest.raises(RuntimeError, match="Unsupported image format"):
        vol.integrate(gray, intrinsic_of(intr), I4)                         # capture_frames.py quirk in reverse
    vol.integrate(rgbd, intrinsic_of(intr), np.linalg.inv(pose))            # 64 units is too small for this frame
    with pytest.raises(RuntimeError, match="overflow"):
        vol.stats()
    with pytest.raises(RuntimeError):
        o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, 8)


def test_tsdf_sharded_union_equals_single(o3r, frames, small_kinect):
    """8(e): each rank keeps the units whose key hashes to it; the union over ranks is the single-GPU volume."""
    intr = small_kinect.depth_intr
    K = intrinsic_of(intr)
    world = 4
    vols = [o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, capacity_units=4000,
                                                         rank=r, world=world) for r in range(world)]
    single, ref = run_both(o3r, frames, intr, 0.01, 0.04, 1, 3)
    for d, c, pose in frames[:3]:
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, False)
        for v in vols:
            v.integrate(rgbd, K, np.linalg.inv(pose))
    seen = {}
    for v in vols:
        keys, tsdf, wgt, _ = v.export_units()
        for i, k in enumerate(keys.cpu().numpy().tolist()):
            assert tuple(k) not in seen
            seen[tuple(k)] = (tsdf[i].cpu().numpy(), wgt[i].cpu().numpy())
    assert set(seen) == {tuple(k) for k in ref.keys().tolist()}
    sizes = [v.num_units() for v in vols]
    assert min(sizes) > 0.5 * max(sizes)                                    # hash partition is balanced
    for k, (t, w) in seen.items():
        rt, rw, _ = ref.unit(np.array(k, np.int32))
        assert np.array_equal(t, rt) and np.array_equal(w, rw)


@pytest.mark.parametrize("world", [2, 3, 5])
def test_sharded_extraction_equals_single(o3r, frames, small_kinect, world):
    """8(e): world ranks emulated in one process -- kx all-gather / slab re-partition (all-to-all-v) / extract own slab / merge
    == the single-GPU mesh and cloud, ELEMENT FOR ELEMENT (same vertex order, same triangle order), and no rank ever holds
    the whole volume."""
    from azurekinect3dreconstruction_b200.pipelines import sharded
    intr = small_kinect.depth_intr
    K = intrinsic_of(intr)
    TT = o3d.pipelines.integration.TSDFVolumeColorType
    single = o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, TT.RGB8, capacity_units=4000)
    vols = [o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, TT.RGB8, capacity_units=4000, rank=r, world=world) for r in range(world)]
    for d, c, pose in frames[:3]:
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, False)
        for v in [single] + vols:
            v.integrate(rgbd, K, np.linalg.inv(pose))
    LC = sharded.LocalComm
    exs = [sharded.SlabExtractor(v, r, world) for r, v in enumerate(vols)]
    kx = LC.all_gather_v([e.unit_kx() for e in exs])                                   # collective 1
    routes = [e.export_routes(kx) for e in exs]                                       # local: per destination (keys, unit records)
    keys = LC.all_to_all_v([routes[r][0] for r in range(world)])                      # collective 2
    units = LC.all_to_all_v([routes[r][1] for r in range(world)])
    for r, e in enumerate(exs):
        for buf, u in zip(e.receive_buffers([int(k.shape[0]) for k in keys[r]]), units[r]):
            buf.copy_(u)                                                                # (the real all-to-all writes into these pool views)
    counts = LC.all_gather_v([e.adopt_and_size(keys[r]) for r, e in enumerate(exs)])    # local + collective 3
    total_units = single.num_units()
    assert sum(e.units_held for e in exs) < 2.5 * total_units and (world < 3 or max(e.units_held for e in exs) < total_units)
    # ---- mesh
    parts = [e.fill_mesh(counts) for e in exs]                                         # local
    mesh = sharded.merge_mesh_parts([[p[i] for p in parts] for i in range(len(parts[0]))])   # collective 4 (gather-v)
    ref = single.extract_triangle_mesh()
    assert sum(int(p[0].shape[0]) for p in parts) == len(ref.vertices) and sum(1 for p in parts if p[0].shape[0]) >= 2
    assert np.array_equal(np.asarray(mesh.vertices), np.asarray(ref.vertices))
    assert np.array_equal(np.asarray(mesh.triangles), np.asarray(ref.triangles))
    assert np.array_equal(np.asarray(mesh.vertex_colors), np.asarray(ref.vertex_colors))
    # ---- surface points
    parts = [e.extract_points_part() for e in exs]
    pcd = sharded.merge_point_parts([[p[i] for p in parts] for i in range(len(parts[0]))])
    rp = single.extract_point_cloud()
    assert np.array_equal(np.asarray(pcd.points), np.asarray(rp.points))
    assert np.array_equal(np.asarray(pcd.normals), np.asarray(rp.normals))
    assert np.array_equal(np.asarray(pcd.colors), np.asarray(rp.colors))


def test_peer_frame_exchange_rings_and_stream_memops(frames, small_kinect):
    """The peer-memory host-frame exchange on ONE rank (a world-1 process group): IPC ring allocation, uploads into ring slots,
    stream-ordered counters (cuStreamWriteValue32 / WaitValue32) pacing slot reuse over more chunks than the ring has slots, and
    b2r_tsdf_integrate_frames on the slots -- the volume must equal direct integration.  (The cross-rank form is checked over
    real IPC / NCCL by tools/sharded_check.py at N = 2 and N = 8.)"""
    import socket
    import torch.distributed as dist
    from azurekinect3dreconstruction_b200.pipelines import sharded
    intr = small_kinect.depth_intr
    K = intrinsic_of(intr)
    n = 9
    depth = torch.from_numpy(np.stack([frames[i % len(frames)][0] for i in range(n)]).view(np.int16)).view(torch.uint16)
    color = torch.stack([torch.from_numpy(frames[i % len(frames)][1]) for i in range(n)])
    E = np.stack([np.linalg.inv(frames[i % len(frames)][2]) for i in range(n)])
    TT = o3d.pipelines.integration.TSDFVolumeColorType
    mk = lambda: o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, TT.RGB8, capacity_units=6000)
    direct = mk()
    direct.integrate_raw(depth.cuda(), color.cuda(), K, E)

    def state(v):
        keys, t, w, c = v.export_units()
        o = torch.as_tensor(np.lexsort(keys.cpu().numpy().T[::-1]), device=t.device)
        return keys[o], t[o], w[o], c[o]
    want = state(direct)
    own_group = not dist.is_initialized()
    if own_group:
        sk = socket.socket(); sk.bind(("127.0.0.1", 0)); port = sk.getsockname()[1]; sk.close()
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=0, world_size=1, device_id=torch.device("cuda", torch.cuda.current_device()))
    try:
        dh, ch = depth.view(torch.int16).pin_memory().view(torch.uint16), color.pin_memory()
        for mode in ("pull", "direct"):
            ex = sharded.PeerFrameExchange(tuple(depth.shape[1:]), 3, per=2, device=torch.device("cuda", torch.cuda.current_device()), slots=2, mode=mode)
            for rep in range(2):                                              # 5 chunks through a 2-slot ring, twice (counters carry on)
                v = mk()
                # (integrate_host_frames by-passes the exchange at world 1: drive it directly)
                ex.run(dh, ch, lambda dp, cp, f0, m: v.integrate_frames(dp, cp, 3, K, E[f0:f0 + m], 1000.0, 3.0))
                got = state(v)
                assert all(torch.equal(a, b) for a, b in zip(got, want)), (mode, rep)
            assert ex.seq == 2 * len(sharded.superchunk_plan(n, 1, 2, True)) and ex.bytes_uploaded == 2 * n * depth.shape[1] * depth.shape[2] * 5
            ex.close()
    finally:
        if own_group:
            dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------------
# a10 / a11 extraction
# ------------------------------------------------------------------------------------------------------
def canon(ids, *arrs):
    order = np.lexsort((ids[:, 3], ids[:, 2], ids[:, 1], ids[:, 0]))
    return (ids[order],) + tuple(a[order] if a is not None else None for a in arrs) + (order,)


def test_extract_point_cloud_parity(o3r, frames, small_kinect):
    vol, ref = run_both(o3r, frames, small_kinect.depth_intr, 0.01, 0.04, 1, 4)
    pcd = vol.extract_point_cloud(with_ids=True)
    xyz, rgb, nrm, ids = ref.extract_points()
    g_ids, g_xyz, g_rgb, g_nrm, _ = canon(pcd.edge_ids.cpu().numpy(), np.asarray(pcd.points), np.asarray(pcd.colors), np.asarray(pcd.normals))
    r_ids, r_xyz, r_rgb, r_nrm, _ = canon(ids, xyz, rgb, nrm)
    assert len(g_ids) == len(r_ids) > 5000
    assert np.array_equal(g_ids, r_ids)                                      # same (voxel, axis) set: bit-exact
    assert np.abs(g_xyz - r_xyz).max() <= 1e-4 * 0.01
    assert np.abs(g_rgb - r_rgb).max() < 1e-5
    fin = np.isfinite(r_nrm).all(1)
    assert np.abs(g_nrm[fin] - r_nrm[fin]).max() < 1e-9


def mesh_signature(vtx, tri, ids):
    """Canonical relabelling: vertices sorted by global edge id; triangles rewritten and sorted."""
    order = np.lexsort((ids[:, 3], ids[:, 2], ids[:, 1], ids[:, 0]))
    rank = np.empty(len(order), np.int64)
    rank[order] = np.arange(len(order))
    t = rank[tri]
    # rotate each triangle so that its smallest index comes first (keeps orientation)
    r = t.argmin(1)
    t = np.stack([t[np.arange(len(t)), (r + k) % 3] for k in range(3)], 1)
    t = t[np.lexsort((t[:, 2], t[:, 1], t[:, 0]))]
    return ids[order], vtx[order], t, order


@pytest.mark.parametrize("color_type", [0, 1])
def test_extract_triangle_mesh_parity(o3r, frames, small_kinect, color_type):
    vol, ref = run_both(o3r, frames, small_kinect.depth_intr, 0.01, 0.04, color_type, 4)
    mesh = vol.extract_triangle_mesh(with_ids=True)
    vtx, vcol, tri, ids = ref.extract_mesh()
    g_ids, g_v, g_t, g_o = mesh_signature(np.asarray(mesh.vertices), np.asarray(mesh.triangles).astype(np.int64), mesh.edge_ids.cpu().numpy())
    r_ids, r_v, r_t, r_o = mesh_signature(vtx, tri.astype(np.int64), ids)
    assert len(g_ids) == len(r_ids) > 5000 and len(g_t) == len(r_t) > 10000
    assert np.array_equal(g_ids, r_ids)                                      # vertex set (global edge ids): bit-exact
    assert np.array_equal(g_t, r_t)                                          # triangle connectivity + winding: bit-exact
    assert np.abs(g_v - r_v).max() <= 1e-4 * 0.01                            # <= 1e-4 voxel
    if color_type:
        assert np.abs(np.asarray(mesh.vertex_colors)[g_o] - vcol[r_o]).max() < 1e-5
    mesh.compute_vertex_normals()
    vn_ref = o3r.mesh_vertex_normals(np.asarray(mesh.vertices), np.asarray(mesh.triangles))
    assert np.abs(np.asarray(mesh.vertex_normals) - vn_ref).max() < 1e-9


def test_extract_analytic_sphere_and_empty(o3r):
    from _helpers import sphere_volume
    r, vl = 0.1, 0.01
    ref = sphere_volume(o3r, r, vl, color_type=1)
    vol = o3d.pipelines.integration.ScalableTSDFVolume(vl, 4 * vl, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, capacity_units=256)
    assert vol.extract_triangle_mesh().is_empty() and vol.extract_point_cloud().is_empty()
    keys = ref.keys()
    units = [ref.unit(k) for k in keys]
    vol.import_units(keys, np.stack([u[0] for u in units]), np.stack([u[1] for u in units]),
                     np.stack([u[2].T for u in units]).astype(np.float32))
    mesh = vol.extract_triangle_mesh(with_ids=True)
    vtx, vcol, tri, ids = ref.extract_mesh()
    g = mesh_signature(np.asarray(mesh.vertices), np.asarray(mesh.triangles).astype(np.int64), mesh.edge_ids.cpu().numpy())
    rr = mesh_signature(vtx, tri.astype(np.int64), ids)
    assert np.array_equal(g[0], rr[0]) and np.array_equal(g[2], rr[2])
    assert np.abs(g[1] - rr[1]).max() <= 1e-4 * vl
    V, F = len(g[0]), len(g[2])
    e = np.sort(np.concatenate([g[2][:, [0, 1]], g[2][:, [1, 2]], g[2][:, [2, 0]]]), axis=1)
    assert V - len(np.unique(e, axis=0)) + F == 2
    # negative unit coordinates are part of this volume (sphere centred on the origin)
    assert (keys < 0).any()
    # run-to-run determinism of the raw (unsorted) output order
    mesh2 = vol.extract_triangle_mesh()
    assert np.array_equal(np.asarray(mesh2.triangles), np.asarray(mesh.triangles)) and np.array_equal(np.asarray(mesh2.vertices), np.asarray(mesh.vertices))


# ------------------------------------------------------------------------------------------------------
# a12 odometry / a13 ICP
# ------------------------------------------------------------------------------------------------------
def rot_angle(R):
    return math.acos(max(-1.0, min(1.0, (np.trace(R) - 1) / 2)))


def test_image_kernels_bit_exact(o3r):
    from azurekinect3dreconstruction_b200 import _lib
    rng = np.random.default_rng(5)
    img = rng.random((37, 53)).astype(np.float32)
    img[10, 10] = np.nan
    t = torch.from_numpy(img).cuda()
    lib = _lib.load()
    for kind in (0, 1, 2):
        out = torch.empty_like(t)
        _lib.check(lib.b2r_image_filter(_lib.ptr(t), 53, 37, kind, _lib.ptr(out), _lib.stream_ptr()))
        assert np.array_equal(out.cpu().numpy(), o3r.image_filter(img, kind), equal_nan=True)
    out = torch.empty((18, 26), dtype=torch.float32, device="cuda")
    _lib.check(lib.b2r_image_downsample(_lib.ptr(t), 53, 37, _lib.ptr(out), _lib.stream_ptr()))
    assert np.array_equal(out.cpu().numpy(), o3r.image_downsample(img), equal_nan=True)


@pytest.mark.parametrize("jac", [0, 1])
def test_odometry_parity(o3r, frames, small_kinect, jac):
    intr = small_kinect.depth_intr
    (ds, cs, _), (dt, ct, _) = frames[1], frames[0]
    mk = lambda c, d: o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, True)
    J = o3d.pipelines.odometry.RGBDOdometryJacobianFromColorTerm() if jac == 0 else o3d.pipelines.odometry.RGBDOdometryJacobianFromHybridTerm()
    opt = o3d.pipelines.odometry.OdometryOption(iteration_number_per_pyramid_level=o3d.utility.IntVector([20, 10, 5]), max_depth_diff=0.03,
                                                min_depth=0.0, max_depth=4.0)
    ok, T, info = o3d.pipelines.odometry.compute_rgbd_odometry(mk(cs, ds), mk(ct, dt), intrinsic_of(intr), I4, J, opt)
    ok_r, T_r, info_r, _ = o3r.odometry(o3r.rgb_to_intensity(cs), o3r.depth_to_float(ds), o3r.rgb_to_intensity(ct), o3r.depth_to_float(dt),
                                        K_of(intr), jac_kind=jac)
    assert ok and ok_r
    assert np.abs(T[:3, 3] - T_r[:3, 3]).max() <= 1e-6                       # <= 1e-6 m
    assert rot_angle(T[:3, :3] @ T_r[:3, :3].T) <= 1e-6                      # <= 1e-6 rad
    assert np.abs(info - info_r).max() <= 1e-6 * np.abs(info_r).max()
    assert info[3, 3] == info_r[3, 3]                                        # correspondence count: exact
    # pair-format check (capture_frames.py:133-151 feeds 3-channel colour -> upstream raises)
    bad = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(cs), o3d.geometry.Image(ds), convert_rgb_to_intensity=False)
    with pytest.raises(RuntimeError, match="same in size"):
        o3d.pipelines.odometry.compute_rgbd_odometry(bad, mk(ct, dt), intrinsic_of(intr), I4, J, opt)


def test_odometry_batched_equals_per_pair_and_oracle(o3r, frames, small_kinect):
    """config 4: the consecutive frames of a sequence are independent pairs (odo_init = I): the batched entry runs them through
    one launch sequence; every pair's result is bit-identical to the per-pair call (and so within 1e-6 of the oracle)."""
    intr = small_kinect.depth_intr
    mk = lambda c, d: o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, True)
    imgs = [mk(c, d) for d, c, _ in frames[:6]]
    J = o3d.pipelines.odometry.RGBDOdometryJacobianFromHybridTerm()
    opt = o3d.pipelines.odometry.OdometryOption(iteration_number_per_pyramid_level=o3d.utility.IntVector([20, 10, 5]), max_depth_diff=0.03,
                                                min_depth=0.0, max_depth=4.0)
    single = [o3d.pipelines.odometry.compute_rgbd_odometry(imgs[i + 1], imgs[i], intrinsic_of(intr), I4, J, opt) for i in range(5)]
    for mb in (5, 2):                                                          # one launch sequence, and chunks of 2 + 2 + 1
        batched = o3d.pipelines.odometry.compute_rgbd_odometry_batch(imgs[1:], imgs[:-1], intrinsic_of(intr), None, J, opt, max_batch=mb)
        assert len(batched) == 5
        for (ok, T, info), (ok_b, T_b, info_b) in zip(single, batched):
            assert ok and ok_b and np.array_equal(T, T_b) and np.array_equal(info, info_b)
    inits = [np.eye(4)] * 5
    again = o3d.pipelines.odometry.compute_rgbd_odometry_batch(imgs[1:], imgs[:-1], intrinsic_of(intr), inits, J, opt)
    assert all(np.array_equal(a[1], b[1]) for a, b in zip(again, single))
    (ds, cs, _), (dt, ct, _) = frames[3], frames[2]
    ok_r, T_r, info_r, _ = o3r.odometry(o3r.rgb_to_intensity(cs), o3r.depth_to_float(ds), o3r.rgb_to_intensity(ct), o3r.depth_to_float(dt), K_of(intr), jac_kind=1)
    assert np.abs(batched[2][1][:3, 3] - T_r[:3, 3]).max() <= 1e-6 and rot_angle(batched[2][1][:3, :3] @ T_r[:3, :3].T) <= 1e-6


def test_icp_parity(o3r, frames, small_kinect):
    intr = small_kinect.depth_intr
    (ds, _, ps), (dt, _, pt) = frames[2], frames[0]
    Kt = K_of(intr)
    src, _, _ = o3r.unproject(o3r.depth_to_float(ds), Kt, stride=2)
    tgt, _, _ = o3r.unproject(o3r.depth_to_float(dt), Kt, stride=2)
    # target normals from the organised depth grid (test-side helper; normal estimation is row f1, out of scope).
    # NB radial "towards the camera" normals would leave rotation about the origin unobservable.
    nrm = image_normals(o3r.depth_to_float(dt), Kt, stride=2)
    assert nrm.shape == tgt.shape
    init = np.linalg.inv(pt) @ ps
    init[:3, 3] += [0.004, -0.003, 0.002]
    s, t = o3d.geometry.PointCloud(src), o3d.geometry.PointCloud(tgt)
    t.normals = o3d.utility.Vector3dVector(nrm)
    crit = o3d.pipelines.registration.ICPConvergenceCriteria(max_iteration=40)
    res = o3d.pipelines.registration.registration_icp(s, t, 0.03, init, o3d.pipelines.registration.TransformationEstimationPointToPlane(), crit)
    T_r, fit_r, rmse_r, corr_r, it_r = o3r.icp_point_to_plane(src, tgt, nrm, 0.03, init, max_iter=40)
    assert res.num_iterations == it_r
    assert np.abs(res.transformation[:3, 3] - T_r[:3, 3]).max() <= 1e-6
    assert rot_angle(res.transformation[:3, :3] @ T_r[:3, :3].T) <= 1e-6
    assert abs(res.fitness - fit_r) <= 1e-9 and abs(res.inlier_rmse - rmse_r) <= 1e-9
    cs = np.asarray(res.correspondence_set)
    ref_pairs = np.stack([np.flatnonzero(corr_r >= 0), corr_r[corr_r >= 0]], 1)
    assert np.array_equal(cs, ref_pairs)                                     # nearest-neighbour result: bit-exact
    # TransformationEstimationPointToPoint (Works2CamerasCOLORVisualizer.py:252-257): no normals needed
    bare_t = o3d.geometry.PointCloud(tgt)
    rq = o3d.pipelines.registration.registration_icp(s, bare_t, 0.03, init, o3d.pipelines.registration.TransformationEstimationPointToPoint(), crit)
    Tq, fitq, rmseq, corrq, itq = o3r.icp_point_to_point(src, tgt, 0.03, init, max_iter=40)
    assert rq.num_iterations == itq
    assert np.abs(rq.transformation[:3, 3] - Tq[:3, 3]).max() <= 1e-6 and rot_angle(rq.transformation[:3, :3] @ Tq[:3, :3].T) <= 1e-6
    assert abs(rq.fitness - fitq) <= 1e-9 and abs(rq.inlier_rmse - rmseq) <= 1e-9
    assert np.array_equal(np.asarray(rq.correspondence_set), np.stack([np.flatnonzero(corrq >= 0), corrq[corrq >= 0]], 1))
    with pytest.raises(RuntimeError, match="with_scaling"):
        o3d.pipelines.registration.registration_icp(s, bare_t, 0.03, init, o3d.pipelines.registration.TransformationEstimationPointToPoint(True), crit)
    t2 = o3d.geometry.PointCloud(tgt)
    with pytest.raises(RuntimeError, match="normal"):
        o3d.pipelines.registration.registration_icp(s, t2, 0.03, init, o3d.pipelines.registration.TransformationEstimationPointToPlane())
    with pytest.raises(RuntimeError):
        o3d.pipelines.registration.registration_icp(s, t, 0.0, init, o3d.pipelines.registration.TransformationEstimationPointToPlane())


def test_posegraph_loop_odometry_optimise_reintegrate(frames, small_kinect):
    """f3, the flow of capture_frames.py:118-198: odometry edges (a12's transformation + information) -> pose graph with a
    loop closure -> global_optimization -> volume.reset() + re-integration under the optimised poses (batched path).
    The re-integrated volume must equal, bit for bit, the same poses integrated frame by frame through the Open3D-shaped call."""
    intr = small_kinect.depth_intr
    K = intrinsic_of(intr)
    Rg = o3d.pipelines.registration
    frames = []                                                            # small inter-frame motion (0.4 deg / 4 mm, as in config 4)
    for p in syn.orbit_trajectory(300)[:6]:
        d, c = small_kinect.render_depth_aligned_rgb(p)
        frames.append((u16(d), c.cpu().numpy(), p))
    mk = lambda c, d: o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, True)
    J = o3d.pipelines.odometry.RGBDOdometryJacobianFromHybridTerm()
    opt = o3d.pipelines.odometry.OdometryOption(iteration_number_per_pyramid_level=o3d.utility.IntVector([20, 10, 5]), max_depth_diff=0.03,
                                                min_depth=0.0, max_depth=4.0)
    n = len(frames)
    pg = Rg.PoseGraph()
    pg.nodes.append(Rg.PoseGraphNode(frames[0][2]))                          # anchored on the true first pose
    pose = frames[0][2].copy()
    for i in range(1, n):
        ok, T, info = o3d.pipelines.odometry.compute_rgbd_odometry(mk(frames[i][1], frames[i][0]), mk(frames[i - 1][1], frames[i - 1][0]), K, I4, J, opt)
        assert ok
        pose = pose @ T                                                    # T maps frame i into frame i-1
        pg.nodes.append(Rg.PoseGraphNode(pose))
        pg.edges.append(Rg.PoseGraphEdge(i - 1, i, np.linalg.inv(T), info, uncertain=False))
    # loop closure: the true relative pose (a perfect re-localisation), weighted like the last odometry edge, uncertain
    pg.edges.append(Rg.PoseGraphEdge(0, n - 1, np.linalg.inv(frames[n - 1][2]) @ frames[0][2], info, uncertain=True))
    drift = max(np.abs(nd.pose - f[2]).max() for nd, f in zip(pg.nodes, frames))
    Rg.global_optimization(pg, Rg.GlobalOptimizationLevenbergMarquardt(), Rg.GlobalOptimizationConvergenceCriteria(),
                           Rg.GlobalOptimizationOption(max_correspondence_distance=0.02, edge_prune_threshold=0.25, preference_loop_closure=1.0, reference_node=0))
    assert pg.last_report["valid"] and len(pg.edges) in (n - 1, n)           # the line process may prune the closure (coarse small-grid odometry)
    assert pg.last_report["residual_after"] <= pg.last_report["residual_before"]
    err = max(np.abs(nd.pose - f[2]).max() for nd, f in zip(pg.nodes, frames))
    assert err < 2e-2 and err <= drift + 1e-6                                # never worse than the chained odometry
    depth = torch.from_numpy(np.stack([f[0] for f in frames]).view(np.int16)).view(torch.uint16).cuda()
    rgb = torch.from_numpy(np.stack([f[1] for f in frames])).cuda()
    TT = o3d.pipelines.integration.TSDFVolumeColorType
    vol = o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, TT.RGB8, capacity_units=4096)
    vol.integrate_raw(depth, rgb, K, np.stack([np.linalg.inv(f[2]) for f in frames]))    # the pre-optimisation volume is discarded
    o3d.pipelines.integration.reintegrate_frames(vol, depth, rgb, pg, K)
    one = o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, TT.RGB8, capacity_units=4096)
    one.set_batch(1)
    for i in range(n):
        img = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(frames[i][1]), o3d.geometry.Image(frames[i][0]), 1000.0, 3.0, False)
        one.integrate(img, K, np.linalg.inv(pg.nodes[i].pose))
    ka, ta, wa, ca = vol.export_units()
    kb, tb, wb, cb = one.export_units()
    oa, ob = np.lexsort(ka.cpu().numpy().T[::-1]), np.lexsort(kb.cpu().numpy().T[::-1])
    assert np.array_equal(ka.cpu().numpy()[oa], kb.cpu().numpy()[ob]) and len(oa) > 50
    ia, ib = torch.as_tensor(oa, device="cuda"), torch.as_tensor(ob, device="cuda")
    assert torch.equal(ta[ia], tb[ib]) and torch.equal(wa[ia], wb[ib]) and torch.equal(ca[ia], cb[ib])


def test_colored_icp_parity(o3r, frames, small_kinect):
    """f4 (part): registration_colored_icp through the C-ABI vs the CPU restatement -- colour gradients of the target,
    correspondences (bit-exact nearest neighbours), pose <= 1e-6 rad / 1e-6 m; upstream's argument errors."""
    intr = small_kinect.depth_intr
    (ds, cs, ps), (dt, ct, pt) = frames[1], frames[0]
    Kt = K_of(intr)
    src, scol, _ = o3r.unproject(o3r.depth_to_float(ds), Kt, rgb=cs, stride=2)
    tgt, tcol, _ = o3r.unproject(o3r.depth_to_float(dt), Kt, rgb=ct, stride=2)
    nrm = image_normals(o3r.depth_to_float(dt), Kt, stride=2)
    init = np.linalg.inv(pt) @ ps
    init[:3, 3] += [0.003, -0.002, 0.002]
    Rg = o3d.pipelines.registration
    s, t = o3d.geometry.PointCloud(src), o3d.geometry.PointCloud(tgt)
    s.colors, t.colors = o3d.utility.Vector3dVector(scol), o3d.utility.Vector3dVector(tcol)
    t.normals = o3d.utility.Vector3dVector(nrm)
    crit = Rg.ICPConvergenceCriteria(relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=40)
    for lam in (0.968, 0.5):
        res = Rg.registration_colored_icp(s, t, 0.03, init, Rg.TransformationEstimationForColoredICP(lambda_geometric=lam), crit)
        T_r, fit_r, rmse_r, corr_r, it_r, grad_r = o3r.icp_colored(src, scol, tgt, nrm, tcol, 0.03, lam, init, max_iter=40)
        g = res.color_gradient.cpu().numpy()
        assert np.abs(g - grad_r).max() <= 1e-9 * max(1.0, np.abs(grad_r).max())
        assert np.abs(grad_r).max() > 1.0                                    # the synthetic checker texture has real gradients
        assert res.num_iterations == it_r
        assert np.abs(res.transformation[:3, 3] - T_r[:3, 3]).max() <= 1e-6
        assert rot_angle(res.transformation[:3, :3] @ T_r[:3, :3].T) <= 1e-6
        assert abs(res.fitness - fit_r) <= 1e-9 and abs(res.inlier_rmse - rmse_r) <= 1e-9
        cs_ = np.asarray(res.correspondence_set)
        assert np.array_equal(cs_, np.stack([np.flatnonzero(corr_r >= 0), corr_r[corr_r >= 0]], 1))
    # the default estimation is lambda 0.968; the result differs from plain point-to-plane (the photometric rows matter)
    plane = Rg.registration_icp(s, t, 0.03, init, Rg.TransformationEstimationPointToPlane(), crit)
    col = Rg.registration_colored_icp(s, t, 0.03, init, criteria=crit)
    assert np.abs(plane.transformation - col.transformation).max() > 1e-7
    bare = o3d.geometry.PointCloud(tgt)
    with pytest.raises(RuntimeError, match="normals"):
        Rg.registration_colored_icp(s, bare, 0.03, init)
    bare.normals = o3d.utility.Vector3dVector(nrm)
    with pytest.raises(RuntimeError, match="target pointcloud to have colors"):
        Rg.registration_colored_icp(s, bare, 0.03, init)
    with pytest.raises(RuntimeError, match="source pointcloud to have colors"):
        Rg.registration_colored_icp(o3d.geometry.PointCloud(src), t, 0.03, init)
    with pytest.raises(RuntimeError, match="lambda_geometric"):
        Rg.registration_colored_icp(s, t, 0.03, init, Rg.TransformationEstimationForColoredICP(lambda_geometric=1.2))
