"""Pins the CPU oracle with analytic known answers: the reference has no tests or golden
vectors and Open3D cannot be installed here, so these closed-form cases are what anchors the oracle.
CPU only (no GPU marker)."""
import math

import numpy as np
import pytest

K = (100.0, 110.0, 31.5, 23.5)
W, H = 64, 48


def plane_depth(D):
    return np.full((H, W), D, np.float32)


def test_depth_conversion_semantics(o3r):
    d = np.array([[0, 999, 1000, 2999, 3000, 65535]], np.uint16)
    f = o3r.depth_to_float(d, 1000.0, 3.0)
    assert f.tolist() == [[0.0, np.float32(999) / np.float32(1000), 1.0, np.float32(2999) / np.float32(1000), 0.0, 0.0]]


def test_intensity_weights(o3r):
    rgb = np.array([[[255, 0, 0], [0, 255, 0], [0, 0, 255], [255, 255, 255]]], np.uint8)
    g = o3r.rgb_to_intensity(rgb)
    np.testing.assert_allclose(g[0], [0.299, 0.587, 0.114, 1.0], rtol=1e-6)


def test_unproject_plane_closed_form(o3r):
    D = 1.25
    xyz, col, idx = o3r.unproject(plane_depth(D), K)
    assert xyz.shape == (W * H, 3)
    assert np.array_equal(idx, np.arange(W * H, dtype=np.int32))          # raster order, all valid
    v, u = np.divmod(np.arange(W * H), W)
    np.testing.assert_allclose(xyz[:, 0], (u - K[2]) * D / K[0], rtol=0, atol=1e-15)
    np.testing.assert_allclose(xyz[:, 1], (v - K[3]) * D / K[1], rtol=0, atol=1e-15)
    assert np.all(xyz[:, 2] == D)


def test_unproject_compaction_and_extrinsic(o3r):
    rng = np.random.default_rng(0)
    d = plane_depth(2.0)
    d[rng.random((H, W)) < 0.3] = 0.0
    rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    th = 0.3
    E = np.eye(4)
    E[:3, :3] = [[math.cos(th), 0, math.sin(th)], [0, 1, 0], [-math.sin(th), 0, math.cos(th)]]
    E[:3, 3] = [0.1, -0.2, 0.3]
    xyz, col, idx = o3r.unproject(d, K, E, rgb=rgb)
    valid = np.flatnonzero(d.reshape(-1) > 0)
    assert np.array_equal(idx, valid.astype(np.int32))
    v, u = np.divmod(valid, W)
    cam = np.stack([(u - K[2]) * 2.0 / K[0], (v - K[3]) * 2.0 / K[1], np.full(valid.size, 2.0), np.ones(valid.size)], 1)
    ref = (np.linalg.inv(E) @ cam.T).T[:, :3]
    np.testing.assert_allclose(xyz, ref, atol=1e-13)
    np.testing.assert_array_equal(col, rgb.reshape(-1, 3)[valid] / 255.0)
    # stride + keep-invalid variants
    xyz4, _, idx4 = o3r.unproject(d, K, stride=4, valid_only=False)
    assert xyz4.shape[0] == (W // 4) * (H // 4)
    assert np.isnan(xyz4[d[::4, ::4].reshape(-1) == 0]).all()
    # empty image
    xyz0, _, idx0 = o3r.unproject(np.zeros((H, W), np.float32), K)
    assert xyz0.shape == (0, 3) and idx0.size == 0


def test_transform_homogeneous(o3r):
    p = np.array([[1.0, 2.0, 3.0]])
    T = np.diag([1.0, -1.0, -1.0, 1.0])
    np.testing.assert_array_equal(o3r.transform_points(p, T), [[1.0, -2.0, -3.0]])
    T2 = np.eye(4)
    T2[3, 3] = 2.0
    np.testing.assert_array_equal(o3r.transform_points(p, T2), [[0.5, 1.0, 1.5]])


def test_block_keys_single_pixel(o3r):
    vl, trunc = 0.004, 0.02
    ul = vl * 16
    d = np.zeros((H, W), np.float32)
    d[8, 12] = 1.0            # a stride-4 sample position
    d[9, 13] = 1.0            # not sampled
    vol = o3r.TSDF(vl, trunc, 0)
    vol.integrate(d, K, np.eye(4))
    p = np.array([(12 - K[2]) * 1.0 / K[0], (8 - K[3]) * 1.0 / K[1], 1.0])
    lo, hi = np.floor((p - trunc) / ul).astype(int), np.floor((p + trunc) / ul).astype(int)
    want = {(a, b, c) for a in range(lo[0], hi[0] + 1) for b in range(lo[1], hi[1] + 1) for c in range(lo[2], hi[2] + 1)}
    assert {tuple(k) for k in vol.keys().tolist()} == want and 1 <= len(want) <= 8


def test_tsdf_plane_closed_form(o3r):
    """One frame of a fronto-parallel plane at depth D, identity pose: every updated voxel holds
    tsdf = min(1, (D - z) * mult / trunc), weight 1; n identical frames leave tsdf unchanged, weight n."""
    vl, trunc, D = 0.01, 0.04, 1.0
    vol = o3r.TSDF(vl, trunc, 0)
    d = plane_depth(D)
    vol.integrate(d, K, np.eye(4))
    keys = vol.keys()
    assert len(keys) > 0
    checked = 0
    for key in keys:
        t, w, _ = vol.unit(key)
        t, w = t.reshape(16, 16, 16), w.reshape(16, 16, 16)
        ii = np.arange(16)
        x = (np.float32(0.005) + np.float32(vl) * ii.astype(np.float32)).astype(np.float64) + key[0] * vl * 16
        y = (np.float32(0.005) + np.float32(vl) * ii.astype(np.float32)).astype(np.float64) + key[1] * vl * 16
        z = 0.005 + vl * ii + key[2] * vl * 16
        X, Y, Z = np.meshgrid(x, y, z, indexing="ij")
        uf = X * K[0] / Z + K[2] + 0.5
        vf = Y * K[1] / Z + K[3] + 0.5
        inside = (Z > 0) & (uf >= 1e-4) & (uf < W - 1e-4) & (vf >= 1e-4) & (vf < H - 1e-4)
        u, v = np.where(inside, uf, 0).astype(int), np.where(inside, vf, 0).astype(int)
        mult = np.sqrt(((u - K[2]) / K[0]) ** 2 + ((v - K[3]) / K[1]) ** 2 + 1)
        sdf = (D - Z) * mult
        upd = inside & (sdf > -trunc)
        # keep clear of decision boundaries where f32 vs f64 rounding may differ
        safe = (np.abs(uf - np.round(uf)) > 1e-3) & (np.abs(vf - np.round(vf)) > 1e-3) & (np.abs(sdf + trunc) > 1e-5)
        ref = np.minimum(1.0, sdf / trunc)
        m = upd & safe
        assert np.all(w[m] == 1.0)
        np.testing.assert_allclose(t[m], ref[m], atol=2e-5)
        assert np.all(w[~upd & safe] == 0.0)
        checked += int(m.sum())
    assert checked > 1000
    before = {tuple(k): vol.unit(k) for k in keys}
    vol.integrate(d, K, np.eye(4))
    vol.integrate(d, K, np.eye(4))
    for k, (t0, w0, _) in before.items():
        t1, w1, _ = vol.unit(k)
        assert np.array_equal(w1, w0 * 3.0)
        np.testing.assert_allclose(t1, t0, atol=1e-6)


def test_tsdf_color_average(o3r):
    vol = o3r.TSDF(0.01, 0.04, 1)
    d = plane_depth(1.0)
    a = np.full((H, W, 3), 100, np.uint8)
    b = np.full((H, W, 3), 200, np.uint8)
    vol.integrate(d, K, np.eye(4), rgb=a)
    vol.integrate(d, K, np.eye(4), rgb=b)
    t, w, c = vol.unit(vol.keys()[0])
    m = w == 2
    assert m.any() and np.allclose(c[m], 150.0)
    with pytest.raises(RuntimeError):
        vol.integrate(d, K, np.eye(4))          # RGB8 volume without colour -> upstream LogError


from _helpers import sphere_volume  # noqa: E402


def test_marching_cubes_sphere(o3r):
    r, vl = 0.1, 0.01
    vol = sphere_volume(o3r, r, vl)
    vtx, _, tri, ids = vol.extract_mesh()
    V, F = len(vtx), len(tri)
    assert V > 500 and F > 1000
    e = np.sort(np.concatenate([tri[:, [0, 1]], tri[:, [1, 2]], tri[:, [2, 0]]]), axis=1)
    ue, cnt = np.unique(e, axis=0, return_counts=True)
    assert np.all(cnt == 2)                         # closed 2-manifold: every edge shared by 2 triangles
    assert V - len(ue) + F == 2                     # Euler characteristic of a sphere
    assert len(np.unique(ids, axis=0)) == V         # welded: one vertex per global edge id
    rad = np.linalg.norm(vtx, axis=1)
    assert np.abs(rad - r).max() < 0.15 * vl        # linear interpolation of a (nearly) linear field
    a, b, c = vtx[tri[:, 0]], vtx[tri[:, 1]], vtx[tri[:, 2]]
    n = np.cross(b - a, c - a)
    area = 0.5 * np.linalg.norm(n, axis=1).sum()
    assert abs(area - 4 * math.pi * r * r) / (4 * math.pi * r * r) < 0.02
    # upstream winding (i, i+2, i+1): for an inside-negative SDF the triangle normals point OUTWARD
    cen = (a + b + c) / 3
    assert np.all((n * cen).sum(1) > 0)
    vn = o3r.mesh_vertex_normals(vtx, tri)
    assert np.all((vn * vtx).sum(1) / rad > 0.95)


def test_surface_points_sphere(o3r):
    r, vl = 0.1, 0.01
    vol = sphere_volume(o3r, r, vl, color_type=1)
    xyz, rgb, nrm, ids = vol.extract_points()
    assert len(xyz) > 500
    rad = np.linalg.norm(xyz, axis=1)
    assert np.abs(rad - r).max() < 0.15 * vl
    assert np.all((nrm * xyz).sum(1) / rad > 0.97)   # gradient of |p| - r is radial
    np.testing.assert_allclose(rgb[:, 0], 0.25, atol=1e-6)
    np.testing.assert_allclose(rgb[:, 1], (128 + 100 * xyz[:, 0] / r) / 255.0, atol=2e-3)
    # every point has a unique (voxel, axis) id and lies on that voxel edge
    assert len(np.unique(ids, axis=0)) == len(ids)
    p0 = 0.5 * vl + vl * ids[:, :3]
    off = xyz - p0
    ax = ids[:, 3]
    along = off[np.arange(len(ids)), ax]
    assert np.all((along >= -1e-12) & (along <= vl + 1e-12))
    off[np.arange(len(ids)), ax] = 0
    assert np.abs(off).max() < 1e-12


def test_missing_neighbour_units_block_cubes(o3r):
    """A single allocated unit: cubes that need a corner from a missing +x/+y/+z unit are skipped."""
    vl = 0.01
    vol = o3r.TSDF(vl, 0.04, 0)
    i = np.arange(16)
    X, Y, Z = np.meshgrid(i, i, i, indexing="ij")
    t = np.clip((X - 7.5) / 4.0, -1, 1).astype(np.float32)       # plane between x=7 and x=8
    vol.import_unit((0, 0, 0), t.reshape(-1), np.ones(4096, np.float32))
    vtx, _, tri, ids = vol.extract_mesh()
    assert len(tri) == 2 * 15 * 15                                   # cubes at y=15 or z=15 are incomplete
    assert np.allclose(vtx[:, 0], 0.5 * vl + vl * 7.5)
    xyz, _, _, pid = vol.extract_points()
    assert len(xyz) == 16 * 16 and np.all(pid[:, 3] == 0)


def test_linear_algebra_helpers(o3r):
    rng = np.random.default_rng(1)
    for _ in range(20):
        M = np.eye(4)
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        M[:3, :3] = q
        M[:3, 3] = rng.normal(size=3)
        np.testing.assert_allclose(o3r.m4_inverse(M), np.linalg.inv(M), atol=1e-13)
        J = rng.normal(size=(50, 6))
        A, b = J.T @ J, rng.normal(size=6)
        np.testing.assert_allclose(o3r.solve6(A, b), np.linalg.solve(A, b), rtol=1e-9, atol=1e-12)
    x = np.array([0.1, -0.2, 0.3, 1.0, 2.0, 3.0])
    T = o3r.vec6_to_m4(x)
    ca, sa, cb, sb, cg, sg = math.cos(.1), math.sin(.1), math.cos(-.2), math.sin(-.2), math.cos(.3), math.sin(.3)
    Rx = np.array([[1, 0, 0], [0, ca, -sa], [0, sa, ca]])
    Ry = np.array([[cb, 0, sb], [0, 1, 0], [-sb, 0, cb]])
    Rz = np.array([[cg, -sg, 0], [sg, cg, 0], [0, 0, 1]])
    np.testing.assert_allclose(T[:3, :3], Rz @ Ry @ Rx, atol=1e-15)
    assert T[:3, 3].tolist() == [1.0, 2.0, 3.0]
    assert np.allclose(o3r.solve6(np.zeros((6, 6)), np.zeros(6)), 0)     # zero system -> zero update (Eigen LDLT behaviour)


def test_filters_and_downsample(o3r):
    rng = np.random.default_rng(2)
    img = rng.random((9, 11)).astype(np.float32)
    g = o3r.image_filter(img, 0)
    pad = np.pad(img.astype(np.float64), 1, mode="edge")
    k = np.array([0.25, 0.5, 0.25])
    h = sum(k[i] * pad[:, i:i + 11] for i in range(3))
    ref = sum(k[j] * h[j:j + 9, :] for j in range(3))
    np.testing.assert_allclose(g, ref, atol=1e-6)
    ramp = np.tile(np.arange(11, dtype=np.float32), (9, 1))
    dx, dy = o3r.image_filter(ramp, 1), o3r.image_filter(ramp, 2)
    assert np.all(dx[:, 1:-1] == 8.0) and np.all(dy == 0.0)          # [-1,0,1] x [1,2,1] on a unit ramp
    ds = o3r.image_downsample(img)
    assert ds.shape == (4, 5)
    np.testing.assert_allclose(ds[1, 2], img[2:4, 4:6].mean(), rtol=1e-6)
    nan_img = img.copy()
    nan_img[4, 4] = np.nan
    assert np.isnan(o3r.image_filter(nan_img, 0)[3:6, 3:6]).all()       # NaN spreads to the 3x3 neighbourhood


def render_pair(small_kinect, motion):
    """Two aligned intensity+depth pairs: target = camera 0, source = camera moved by `motion` (so that
    the source->target transform equals `motion`)."""
    import torch
    from azurekinect3dreconstruction_b200 import synthetic as syn
    pose_t = syn.orbit_trajectory(5)[2]
    pose_s = pose_t @ motion
    out = []
    for pose in (pose_s, pose_t):
        dep, rgb = small_kinect.render_depth_aligned_rgb(pose)
        d = dep.view(torch.int16).numpy().view(np.uint16)
        out.append((rgb.numpy(), d))
    return out


@pytest.mark.parametrize("jac", [0, 1])
def test_odometry_recovers_known_motion(o3r, jac):
    from azurekinect3dreconstruction_b200 import synthetic as syn
    kin = syn.SyntheticKinect(depth_intrinsics=dict(width=160, height=144, fx=126.05, fy=126.025, cx=79.95, cy=71.975), seed=0)
    motion = syn.make_pose(syn.rot_y(math.radians(0.5)) @ syn.rot_x(math.radians(-0.3)), (0.01, -0.004, 0.006))
    (rs, ds), (rt, dt) = render_pair(kin, motion)
    Kd = (126.05, 126.025, 79.95, 71.975)
    gs, gt = o3r.rgb_to_intensity(rs), o3r.rgb_to_intensity(rt)
    fs, ft = o3r.depth_to_float(ds, 1000.0, 3.0), o3r.depth_to_float(dt, 1000.0, 3.0)
    ok, T, info, dbg = o3r.odometry(gs, fs, gt, ft, Kd, jac_kind=jac)
    assert ok and dbg[0] > 1000
    dR = T[:3, :3] @ motion[:3, :3].T
    ang = math.degrees(math.acos(min(1.0, (np.trace(dR) - 1) / 2)))
    # 1 mm depth quantisation + 160x144 pixels: expect ~1 mm / 0.1 degree accuracy
    tol_t, tol_a = (0.006, 0.4) if jac == 0 else (0.003, 0.15)
    assert np.linalg.norm(T[:3, 3] - motion[:3, 3]) < tol_t and ang < tol_a
    assert np.allclose(info, info.T) and np.all(np.linalg.eigvalsh(info) > 0)
    assert info[3, 3] == info[4, 4] == info[5, 5] and info[3, 3] > 1000      # = number of correspondences


def test_odometry_identical_pair_is_identity(o3r):
    """bundlefusion_reconstruction.py:224-225 passes the same rgbd as source and target."""
    from azurekinect3dreconstruction_b200 import synthetic as syn
    kin = syn.SyntheticKinect(depth_intrinsics=dict(width=160, height=144, fx=126.05, fy=126.025, cx=79.95, cy=71.975), seed=0)
    (rs, ds), _ = render_pair(kin, np.eye(4))
    g, f = o3r.rgb_to_intensity(rs), o3r.depth_to_float(ds, 1000.0, 3.0)
    ok, T, info, _ = o3r.odometry(g, f, g, f, (126.05, 126.025, 79.95, 71.975))
    assert ok
    np.testing.assert_allclose(T, np.eye(4), atol=1e-9)


def test_hybrid_jacobian_matches_finite_differences(o3r):
    """Gauss-Newton step direction check: one iteration from identity reduces the true pose error."""
    from azurekinect3dreconstruction_b200 import synthetic as syn
    kin = syn.SyntheticKinect(depth_intrinsics=dict(width=160, height=144, fx=126.05, fy=126.025, cx=79.95, cy=71.975), seed=0)
    motion = syn.make_pose(syn.rot_y(math.radians(0.3)), (0.006, 0.0, 0.0))
    (rs, ds), (rt, dt) = render_pair(kin, motion)
    Kd = (126.05, 126.025, 79.95, 71.975)
    a = (o3r.rgb_to_intensity(rs), o3r.depth_to_float(ds), o3r.rgb_to_intensity(rt), o3r.depth_to_float(dt))
    errs = []
    for n in (1, 3, 8):
        ok, T, _, _ = o3r.odometry(*a, Kd, jac_kind=1, iters=(n,))
        errs.append(np.linalg.norm(T[:3, 3] - motion[:3, 3]))
    assert errs[0] < 0.006 and errs[2] <= errs[0]


def test_icp_recovers_inverse_motion(o3r):
    rng = np.random.default_rng(3)
    # target: points on a bumpy surface with exact normals
    u, v = rng.random(4000) * 2 - 1, rng.random(4000) * 2 - 1
    z = 0.2 * np.sin(2 * u) * np.cos(1.5 * v) + 0.1 * u * v
    tgt = np.stack([u, v, z], 1)
    nz = np.stack([-(0.4 * np.cos(2 * u) * np.cos(1.5 * v) + 0.1 * v), -(-0.3 * np.sin(2 * u) * np.sin(1.5 * v) + 0.1 * u), np.ones_like(u)], 1)
    nrm = nz / np.linalg.norm(nz, axis=1, keepdims=True)
    from azurekinect3dreconstruction_b200 import synthetic as syn
    M = syn.make_pose(syn.rot_z(math.radians(1.0)) @ syn.rot_x(math.radians(0.7)), (0.01, -0.008, 0.012))
    src = o3r.transform_points(tgt[::2], M)
    T, fit, rmse, corr, it = o3r.icp_point_to_plane(src, tgt, nrm, 0.05, max_iter=60)
    assert fit > 0.99 and rmse < 2e-3 and 1 <= it <= 60
    np.testing.assert_allclose(T @ M, np.eye(4), atol=2e-3)
    assert (corr >= 0).mean() > 0.99
    # exact-NN check against brute force at the final pose
    fin = o3r.transform_points(src, T)
    d2 = ((fin[:200, None, :] - tgt[None, :, :]) ** 2).sum(-1)
    assert np.array_equal(corr[:200], d2.argmin(1).astype(np.int32))
    with pytest.raises(RuntimeError):
        o3r.icp_point_to_plane(src, tgt, nrm, 0.0)
    with pytest.raises(RuntimeError):
        o3r.icp_point_to_plane(src, tgt, None, 0.05)
    # TransformationEstimationPointToPoint: same loop, closed-form rigid fit per iteration (slower convergence, same fixed point)
    Tq, fitq, rmseq, corrq, itq = o3r.icp_point_to_point(src, tgt, 0.05, max_iter=200)
    assert fitq > 0.99 and 1 <= itq <= 200
    np.testing.assert_allclose(Tq @ M, np.eye(4), atol=5e-3)
    np.testing.assert_allclose(Tq[:3, :3] @ Tq[:3, :3].T, np.eye(3), atol=1e-12)
    with pytest.raises(RuntimeError):
        o3r.icp_point_to_point(src, tgt, -1.0)
    # nothing within reach: identity update, fitness 0
    T0, fit0, rmse0, _, _ = o3r.icp_point_to_plane(src + 100.0, tgt, nrm, 0.05, max_iter=5)
    assert fit0 == 0.0 and rmse0 == 0.0 and np.array_equal(T0, np.eye(4))


def test_colored_icp_gradient_is_analytic_and_colour_locks_the_in_plane_slide(o3r):
    """f4 (registration_colored_icp): on a plane carrying a linear intensity the tangent-plane gradient is the analytic one;
    a slide WITHIN the plane is invisible to point-to-plane ICP (zero residual) and recovered by the photometric term."""
    g = np.arange(-30, 31) * 0.01
    X, Y = np.meshgrid(g, g, indexing="ij")
    rng = np.random.default_rng(0)
    tgt = np.stack([X.ravel() + rng.normal(0, 1e-4, X.size), Y.ravel() + rng.normal(0, 1e-4, X.size), np.full(X.size, 2.0)], 1)
    nrm = np.tile([0.0, 0.0, 1.0], (len(tgt), 1))
    inside = (np.abs(tgt[:, 0]) < 0.25) & (np.abs(tgt[:, 1]) < 0.25)
    # (1) linear intensity: exact gradient (max_iter = 0 only builds the gradients and evaluates)
    a, b = 0.8, -0.5
    lin = np.repeat((0.5 + a * tgt[:, 0] + b * tgt[:, 1])[:, None], 3, 1)
    _, fit0, _, _, it0, grad = o3r.icp_colored(tgt[inside], lin[inside], tgt, nrm, lin, 0.02, max_iter=0)
    assert it0 == 0 and fit0 == 1.0
    np.testing.assert_allclose(grad[inside], np.tile([a, b, 0.0], (inside.sum(), 1)), atol=1e-9)
    # (2) an intensity that varies in both directions makes both in-plane translations observable
    inten = lambda p: 0.5 + 0.25 * np.sin(9.0 * p[:, 0]) + 0.25 * np.cos(7.0 * p[:, 1] + 0.3)
    tcol = np.repeat(inten(tgt)[:, None], 3, 1)
    shift = np.array([0.004, -0.003, 0.0])
    inner = (np.abs(tgt[:, 0]) < 0.2) & (np.abs(tgt[:, 1]) < 0.2)
    src = tgt[inner] + rng.normal(0, 2e-4, (inner.sum(), 3)) * [1, 1, 0]
    scol = np.repeat(inten(src)[:, None], 3, 1)                              # the colour the surface has where the source point IS
    src_shifted = src - shift                                               # expressed in a frame displaced by +shift
    T, fit, rmse, corr, it, grad = o3r.icp_colored(src_shifted, scol, tgt, nrm, tcol, 0.02, max_iter=50)
    assert fit == 1.0 and 1 <= it <= 50
    np.testing.assert_allclose(T[:3, 3], shift, atol=3e-4)                   # finite-difference gradients over a 4 cm neighbourhood
    np.testing.assert_allclose(T[:3, :3], np.eye(3), atol=2e-3)
    Tp, *_ = o3r.icp_point_to_plane(src_shifted, tgt, nrm, 0.02, max_iter=50)
    assert np.abs(Tp[:2, 3]).max() < 1e-9                                    # geometry alone cannot see the slide
    for bad in (dict(tgt_nrm=None), dict(tgt_rgb=None), dict(src_rgb=None)):
        kw = dict(src_rgb=scol, tgt_nrm=nrm, tgt_rgb=tcol); kw.update(bad)
        with pytest.raises(RuntimeError):
            o3r.icp_colored(src_shifted, kw["src_rgb"], tgt, kw["tgt_nrm"], kw["tgt_rgb"], 0.02)
    with pytest.raises(RuntimeError):
        o3r.icp_colored(src_shifted, scol, tgt, nrm, tcol, 0.02, lambda_geometric=1.5)


def test_fpfh_oracle_known_answers(o3r):
    """f4 (compute_fpfh_feature): each of the three 11-bin blocks sums to 200 (100 from the weighted neighbour SPFHs + 100 from
    the point's own SPFH) wherever a point has neighbours; on a PLANE with parallel normals every pair has alpha = phi = 0 and
    theta = 0, so all mass sits in the middle bin of each block; the descriptor is invariant under rigid motion."""
    rng = np.random.default_rng(1)
    g = np.arange(-10, 11) * 0.01
    X, Y = np.meshgrid(g, g, indexing="ij")
    plane = np.stack([X.ravel(), Y.ravel(), np.zeros(X.size)], 1) + rng.normal(0, 1e-5, (X.size, 3)) * [1, 1, 0]
    nrm = np.tile([0.0, 0.0, 1.0], (len(plane), 1))
    f = o3r.compute_fpfh(plane, nrm, max_nn=30, radius=0.035)
    assert f.shape == (len(plane), 33)
    np.testing.assert_allclose(f.reshape(-1, 3, 11).sum(2), 200.0, atol=1e-9)
    assert np.all(f[:, 5] > 199.999) and np.all(f[:, 16] > 199.999) and np.all(f[:, 27] > 199.999)
    # a curved, bumpy surface: sums still 200, and rigid motion leaves the descriptor unchanged up to rounding
    u, v = rng.random(1500) * 0.4 - 0.2, rng.random(1500) * 0.4 - 0.2
    z = 0.05 * np.sin(12 * u) * np.cos(9 * v)
    pts = np.stack([u, v, z], 1)
    nz = np.stack([-0.6 * np.cos(12 * u) * np.cos(9 * v), 0.45 * np.sin(12 * u) * np.sin(9 * v), np.ones_like(u)], 1)
    nn = nz / np.linalg.norm(nz, axis=1, keepdims=True)
    f1 = o3r.compute_fpfh(pts, nn, max_nn=50, radius=0.05)
    has = f1.sum(1) > 0
    assert has.mean() > 0.99
    np.testing.assert_allclose(f1[has].reshape(-1, 3, 11).sum(2), 200.0, atol=1e-8)
    assert (f1[has] > 0).sum(1).mean() > 4.5                                 # spread over more bins than the plane (exactly 3)
    from azurekinect3dreconstruction_b200 import synthetic as syn
    M = syn.make_pose(syn.rot_z(0.7) @ syn.rot_x(-0.4), (0.3, -0.2, 1.1))
    f2 = o3r.compute_fpfh(o3r.transform_points(pts, M), nn @ M[:3, :3].T, max_nn=50, radius=0.05)
    assert np.abs(f1 - f2).max() < 25.0 and np.median(np.abs(f1 - f2)) < 1e-9  # only bin-edge flips / distance ties may move mass
    with pytest.raises(RuntimeError):
        o3r.compute_fpfh(pts, None)


def test_ransac_oracle_known_answers(o3r):
    """f4 (registration_ransac_based_on_*): Horn's closed form equals the SVD (Kabsch / umeyama without scale) optimum; with
    60 % wrong correspondences RANSAC still recovers the motion; checkers reject inconsistent samples; degenerate arguments
    return upstream's empty result; same seed -> same answer."""
    from azurekinect3dreconstruction_b200 import synthetic as syn
    rng = np.random.default_rng(7)
    for n in (3, 4, 8):
        s = rng.normal(0, 1, (n, 3))
        M = syn.make_pose(syn.rot_z(0.9) @ syn.rot_y(-0.5) @ syn.rot_x(0.3), (0.4, -1.2, 0.7))
        t = o3r.transform_points(s, M) + rng.normal(0, 1e-3, (n, 3))
        T = o3r.horn_rigid(s, t)
        ms, mt = s.mean(0), t.mean(0)
        U, _, Vt = np.linalg.svd((t - mt).T @ (s - ms))
        D = np.diag([1.0, 1.0, np.sign(np.linalg.det(U @ Vt))])
        Rk = U @ D @ Vt
        np.testing.assert_allclose(T[:3, :3], Rk, atol=1e-9)
        np.testing.assert_allclose(T[:3, 3], mt - Rk @ ms, atol=1e-9)
        np.testing.assert_allclose(T[:3, :3] @ T[:3, :3].T, np.eye(3), atol=1e-12)
    # a bumpy surface, its moved copy, and a correspondence list that is 60 % garbage
    u, v = rng.random(600) * 2 - 1, rng.random(600) * 2 - 1
    tgt = np.stack([u, v, 0.3 * np.sin(3 * u) * np.cos(2 * v) + 0.2 * u * v], 1)
    M = syn.make_pose(syn.rot_z(0.6) @ syn.rot_x(-0.35), (0.5, -0.3, 0.8))
    src = o3r.transform_points(tgt, np.linalg.inv(M)) + rng.normal(0, 2e-4, tgt.shape)
    corres = np.stack([np.arange(600), np.arange(600)], 1).astype(np.int32)
    bad = rng.random(600) < 0.6
    corres[bad, 1] = rng.integers(0, 600, bad.sum())
    T, fit, rmse, rep = o3r.ransac_correspondence(src, tgt, corres, 0.02, ransac_n=4, edge_thr=0.9, dist_thr=0.02, max_iteration=20000, confidence=0.999, seed=1)
    np.testing.assert_allclose(T, M, atol=5e-3)
    assert fit > 0.95 and rmse < 5e-3 and 0 <= rep[2] < rep[0] <= 20000 and rep[1] >= 1
    assert rep[0] < 20000                                                    # the confidence rule cut the budget short
    assert rep[3] >= 0.3 * 600                                               # the winner explains the clean 40 %
    T2, fit2, rmse2, rep2 = o3r.ransac_correspondence(src, tgt, corres, 0.02, ransac_n=4, edge_thr=0.9, dist_thr=0.02, max_iteration=20000, confidence=0.999, seed=1)
    assert np.array_equal(T, T2) and rep == rep2                             # deterministic for a seed
    _, _, _, rep3 = o3r.ransac_correspondence(src, tgt, corres, 0.02, ransac_n=4, max_iteration=300, confidence=1.0, seed=1)
    assert rep3[0] == 300 and rep3[1] == 300                                 # no checkers: every draw is validated; confidence 1 never stops early
    Te, fe, re_, repe = o3r.ransac_correspondence(src, tgt, corres[:2], 0.02, ransac_n=3)
    assert np.array_equal(Te, np.eye(4)) and fe == 0.0 and repe[0] == 0      # fewer correspondences than ransac_n: empty result
    Te, fe, _, _ = o3r.ransac_correspondence(src, tgt, corres, 0.0, ransac_n=3)
    assert np.array_equal(Te, np.eye(4)) and fe == 0.0
    # nearest-feature matching: exact argmin, ties to the lower index
    fq, ft = rng.random((50, 33)), rng.random((70, 33))
    ft[10] = ft[3]
    fq[0] = ft[3]
    m = o3r.match_features(fq, ft)
    assert m[0] == 3 and np.array_equal(m, ((fq[:, None, :] - ft[None]) ** 2).sum(-1).argmin(1).astype(np.int32))
