"""GPU parity tests for the point-cloud conditioning row: exact neighbour search, estimate_normals,
orient_normals_towards_camera_location, remove_statistical_outlier, voxel_down_sample -- the CUDA path through the C-ABI
vs the brute-force CPU oracle (oracle/oracle_pcd.cpp) on identical inputs.
Bars: neighbour indices, inlier masks, voxel sets: bit-exact; covariances, mean distances, voxel averages: bit-exact f64
(same summation order); normals: |n . n_ref| >= 1 - 1e-9 (the eigen solver calls acos / cos, whose last bit differs
between glibc and CUDA)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import azurekinect3dreconstruction_b200 as o3d
from azurekinect3dreconstruction_b200 import _lib, synthetic as syn
from azurekinect3dreconstruction_b200.geometry import _Knn


def intrinsic_of(intr):
    return o3d.camera.PinholeCameraIntrinsic(intr["width"], intr["height"], intr["fx"], intr["fy"], intr["cx"], intr["cy"])


@pytest.fixture(scope="module")
def cloud(small_kinect):
    """One unprojected synthetic frame (about 22 k coloured points, a surface sampled as a depth camera samples it)."""
    pose = syn.orbit_trajectory(4, yaw_sweep_deg=10.0)[2]
    d, c = small_kinect.render_depth_aligned_rgb(pose)
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, False)
    return o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intrinsic_of(small_kinect.depth_intr))


def dev(a, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).cuda().contiguous()


def gpu_search(xyz, k, radius, cell, query=None):
    t = dev(xyz)
    knn = _Knn(t, cell, k)
    nq = len(xyz) if query is None else len(query)
    q = None if query is None else dev(query)
    idx = torch.empty((nq, k), dtype=torch.int32, device="cuda")
    d2 = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    cnt = torch.empty((nq,), dtype=torch.int32, device="cuda")
    _lib.check(_lib.load().b2r_knn_search(knn._h, _lib.ptr(q), nq, k, float(radius), _lib.ptr(idx), _lib.ptr(d2), _lib.ptr(cnt), _lib.stream_ptr()))
    torch.cuda.synchronize()
    knn.close()
    return idx.cpu().numpy(), d2.cpu().numpy(), cnt.cpu().numpy()


@pytest.mark.parametrize("k,radius,cell", [(20, -1.0, -1.0), (30, 0.04, 0.04), (30, 0.1, 0.1), (8, -1.0, 0.5), (40, 0.05, 0.02)])
def test_knn_search_exact(o3r, cloud, k, radius, cell):
    xyz = np.asarray(cloud.points)
    idx, d2, cnt = gpu_search(xyz, k, radius, cell)
    ridx, rd2, rcnt = o3r.pcd_search(xyz, None, k, radius)
    assert np.array_equal(cnt, rcnt)
    assert np.array_equal(idx, ridx)                       # neighbour sets AND order: bit-exact
    assert np.array_equal(d2, rd2)
    assert (idx[:, 0] == np.arange(len(xyz))).all()        # a point is its own first neighbour


def test_knn_random_cloud_duplicates_and_external_queries(o3r):
    rng = np.random.default_rng(5)
    pts = rng.uniform(-1, 1, (6000, 3))
    pts[100:110] = pts[100]                                # exact duplicates: ties resolved by index
    pts = np.r_[pts, [[5.0, 5.0, 5.0]]]                    # an isolated point far away (forces many rings)
    q = rng.uniform(-1.5, 1.5, (500, 3))                   # queries outside the grid too
    for k, r in ((12, -1.0), (16, 0.3)):
        idx, d2, cnt = gpu_search(pts, k, r, -1.0 if r < 0 else r, query=q)
        ridx, rd2, rcnt = o3r.pcd_search(pts, q, k, r)
        assert np.array_equal(cnt, rcnt) and np.array_equal(idx, ridx) and np.array_equal(d2, rd2)
    idx, _, _ = gpu_search(pts, 12, -1.0, -1.0)
    ridx, _, _ = o3r.pcd_search(pts, None, 12, -1.0)
    assert np.array_equal(idx, ridx)
    # fewer points than k
    few = pts[:5]
    idx, _, cnt = gpu_search(few, 8, -1.0, -1.0)
    assert (cnt == 5).all() and (idx[:, 5:] == -1).all()


@pytest.mark.parametrize("k,radius", [(30, 0.04), (30, -1.0), (30, 0.1)])
def test_estimate_normals_parity(o3r, cloud, k, radius):
    xyz = np.asarray(cloud.points)
    t = dev(xyz)
    knn = _Knn(t, radius if radius > 0 else -1.0, k)
    nrm = torch.zeros_like(t)
    cov = torch.empty((len(xyz), 6), dtype=torch.float64, device="cuda")
    _lib.check(_lib.load().b2r_pcd_estimate_normals(knn._h, _lib.ptr(t), k, float(radius), _lib.ptr(nrm), 0, _lib.ptr(cov), _lib.stream_ptr()))
    torch.cuda.synchronize()
    knn.close()
    rn, rcov = o3r.pcd_estimate_normals(xyz, k, radius)
    assert np.array_equal(cov.cpu().numpy(), rcov)         # same neighbours, same summation order: bit-exact f64
    n = nrm.cpu().numpy()
    dots = np.einsum("ij,ij->i", n, rn)
    assert np.abs(np.abs(dots) - 1.0).max() <= 1e-9
    assert (dots > 0).mean() > 0.999                       # same branch of the solver, same sign
    assert np.abs(np.linalg.norm(n, axis=1) - 1.0).max() < 1e-12
    # the public call (what the reference scripts do), then orientation towards the camera: exact given equal inputs
    pcd = o3d.geometry.PointCloud(t.clone())
    pcd.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius, k) if radius > 0 else o3d.geometry.KDTreeSearchParamKNN(k))
    assert np.array_equal(np.asarray(pcd.normals), n)
    pcd.orient_normals_towards_camera_location(np.array([0.0, 0.0, 0.0]))
    assert np.array_equal(np.asarray(pcd.normals), o3r.pcd_orient_normals(xyz, n))
    assert (np.einsum("ij,ij->i", np.asarray(pcd.normals), -xyz) >= 0).all()
    # a second estimate keeps the orientation of the existing normals (has_normals path)
    before = np.asarray(pcd.normals).copy()
    pcd.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius, k) if radius > 0 else o3d.geometry.KDTreeSearchParamKNN(k))
    assert (np.einsum("ij,ij->i", np.asarray(pcd.normals), before) > 0).all()


def test_estimate_normals_plane_known_answer():
    rng = np.random.default_rng(2)
    xy = rng.uniform(-1, 1, (5000, 2))
    pts = np.c_[xy, 0.3 * xy[:, 0] + 0.1 * xy[:, 1] + 1.0]
    pcd = o3d.geometry.PointCloud(dev(pts))
    pcd.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(0.2, 30))
    ref = np.array([0.3, 0.1, -1.0]) / np.linalg.norm([0.3, 0.1, -1.0])
    assert np.abs(np.abs(np.asarray(pcd.normals) @ ref) - 1.0).max() < 1e-9
    lonely = o3d.geometry.PointCloud(dev([[0.0, 0, 0], [10.0, 0, 0]]))      # fewer than 3 neighbours -> (0, 0, 1)
    lonely.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(0.1, 30))
    assert np.array_equal(np.asarray(lonely.normals), [[0, 0, 1.0], [0, 0, 1.0]])


def test_remove_statistical_outlier_parity(o3r, cloud):
    xyz = np.asarray(cloud.points)
    rng = np.random.default_rng(9)
    xyz = np.r_[xyz, rng.uniform(-2, 2, (40, 3)) + [0, 0, 2.0]]          # floating outliers
    pcd = o3d.geometry.PointCloud(dev(xyz))
    pcd.colors = o3d.utility.Vector3dVector(dev(rng.uniform(0, 1, xyz.shape)))
    t = pcd.points.tensor
    knn = _Knn(t, -1.0, 20)
    keep = torch.empty((len(xyz),), dtype=torch.uint8, device="cuda")
    idx = torch.empty((len(xyz),), dtype=torch.int32, device="cuda")
    avg = torch.empty((len(xyz),), dtype=torch.float64, device="cuda")
    m = C.c_int64(0)
    _lib.check(_lib.load().b2r_pcd_statistical_outlier(knn._h, 20, 2.0, _lib.ptr(keep), _lib.ptr(idx), C.byref(m), _lib.ptr(avg), _lib.stream_ptr()))
    knn.close()
    rkeep, ravg = o3r.pcd_statistical_outlier(xyz, 20, 2.0)
    assert np.array_equal(avg.cpu().numpy(), ravg)                         # mean neighbour distance: bit-exact
    assert np.array_equal(keep.cpu().numpy().astype(bool), rkeep)          # inlier mask: bit-exact
    assert m.value == int(rkeep.sum()) and np.array_equal(idx[:m.value].cpu().numpy(), np.nonzero(rkeep)[0])
    assert rkeep[-40:].sum() < 10 and rkeep[:-40].mean() > 0.9
    inl, ind = pcd.remove_statistical_outlier(nb_neighbors=20, std_ratio=2.0)                # the reference's call
    assert np.array_equal(np.asarray(inl.points), xyz[rkeep]) and np.array_equal(np.asarray(ind.cpu()), np.nonzero(rkeep)[0])
    assert np.array_equal(np.asarray(inl.colors), np.asarray(pcd.colors)[rkeep])
    with pytest.raises(RuntimeError):
        pcd.remove_statistical_outlier(0, 2.0)
    with pytest.raises(RuntimeError):
        pcd.remove_statistical_outlier(20, 0.0)


@pytest.mark.parametrize("voxel", [0.01, 0.02, 0.25])
def test_voxel_down_sample_parity(o3r, cloud, voxel):
    xyz, col = np.asarray(cloud.points), np.asarray(cloud.colors)
    nrm, _ = o3r.pcd_estimate_normals(xyz[:2000], 10, -1.0)
    nrm = np.r_[nrm, np.tile([[0.0, 0.0, 1.0]], (len(xyz) - 2000, 1))]
    nrm[5] = np.nan                                                        # upstream skips NaN normals in the sum
    pcd = o3d.geometry.PointCloud(dev(xyz))
    pcd.colors = o3d.utility.Vector3dVector(dev(col))
    pcd.normals = o3d.utility.Vector3dVector(dev(nrm))
    down = pcd.voxel_down_sample(voxel)
    rp, rc, rn, rk = o3r.pcd_voxel_down_sample(xyz, voxel, col, nrm)
    assert len(down.points) == len(rp) < len(xyz)
    assert np.array_equal(down.voxel_index.cpu().numpy(), rk)              # voxel set (and canonical order): bit-exact
    assert np.array_equal(np.asarray(down.points), rp)                     # sequential sums in input order: bit-exact
    assert np.array_equal(np.asarray(down.colors), rc) and np.array_equal(np.asarray(down.normals), rn)
    plain = o3d.geometry.PointCloud(dev(xyz)).voxel_down_sample(voxel)     # no colours / normals
    assert np.array_equal(np.asarray(plain.points), rp) and not plain.has_colors() and not plain.has_normals()
    with pytest.raises(RuntimeError):
        pcd.voxel_down_sample(0.0)
    assert np.array_equal(pcd.get_min_bound(), xyz.min(0)) and np.array_equal(pcd.get_max_bound(), xyz.max(0))


def test_preprocess_then_icp_like_the_reference(o3r, small_kinect):
    """GoodColorReconstruction.py:141-153 + 194-198: outlier removal -> voxel_down_sample -> estimate_normals ->
    orient -> point-to-plane ICP between two consecutive frames recovers the relative motion."""
    poses = syn.orbit_trajectory(30, yaw_sweep_deg=30.0)
    K = intrinsic_of(small_kinect.depth_intr)
    clouds = []
    for p in (poses[0], poses[1]):
        d, c = small_kinect.render_depth_aligned_rgb(p)
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, False)
        pcd = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, K)
        pcd, _ = pcd.remove_statistical_outlier(nb_neighbors=20, std_ratio=2.0)
        pcd = pcd.voxel_down_sample(0.03)
        pcd.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=0.1, max_nn=30))
        pcd.orient_normals_towards_camera_location(np.array([0, 0, 0]))
        clouds.append(pcd)
    src, tgt = clouds[1], clouds[0]
    res = o3d.pipelines.registration.registration_icp(src, tgt, 0.06, np.eye(4), o3d.pipelines.registration.TransformationEstimationPointToPlane(),
                                                      o3d.pipelines.registration.ICPConvergenceCriteria(max_iteration=50))
    gt = np.linalg.inv(poses[0]) @ poses[1]
    assert res.fitness > 0.8
    assert np.linalg.norm(res.transformation[:3, 3] - gt[:3, 3]) < 5e-3
    assert np.abs(res.transformation[:3, :3] - gt[:3, :3]).max() < 5e-3


def test_fpfh_parity(o3r, cloud):
    """f4: compute_fpfh_feature on the down-sampled cloud of preprocess_point_cloud (BetterVisualizer2CamWorksCOLOR.py:197-216).
    Same neighbour order and the same f64 summation order as the oracle: identical histograms except where glibc's and
    CUDA's atan2 differ in the last bit exactly on a bin edge (none expected; a handful tolerated)."""
    down = cloud.voxel_down_sample(0.03)
    down.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=0.06, max_nn=30))
    xyz, nrm = np.asarray(down.points), np.asarray(down.normals)
    assert len(xyz) > 2000
    for param, kw in ((o3d.geometry.KDTreeSearchParamHybrid(radius=0.15, max_nn=100), dict(max_nn=100, radius=0.15)),
                      (o3d.geometry.KDTreeSearchParamKNN(knn=25), dict(max_nn=25, radius=0.0))):
        feat = o3d.pipelines.registration.compute_fpfh_feature(down, param)
        assert feat.dimension() == 33 and feat.num() == len(xyz) and feat.data.shape == (33, len(xyz))
        got = feat.tensor.cpu().numpy()
        want = o3r.compute_fpfh(xyz, nrm, **kw)
        rows_off = (np.abs(got - want).max(1) > 1e-9).sum()
        assert rows_off <= max(2, len(xyz) // 2000), rows_off
        ok = np.abs(got - want).max(1) <= 1e-9
        assert np.array_equal(got[ok], want[ok])                             # where no bin flipped: bit-identical f64
        np.testing.assert_allclose(got.reshape(-1, 3, 11).sum(2)[got.sum(1) > 0], 200.0, atol=1e-8)
    bare = o3d.geometry.PointCloud(xyz)
    with pytest.raises(RuntimeError, match="no normal"):
        o3d.pipelines.registration.compute_fpfh_feature(bare, o3d.geometry.KDTreeSearchParamHybrid(radius=0.15, max_nn=100))


def test_ransac_feature_matching_parity_and_reference_call(o3r, small_kinect):
    """f4: detect_and_estimate_transformation of BetterVisualizer2CamWorksCOLOR.py:219-243 -- two views of the scene,
    preprocess (down-sample, normals, FPFH), nearest-feature correspondences, RANSAC.  The correspondence list is bit-exact
    vs the oracle; the RANSAC winner (same seed, same draws, same rigid fits) is the same iteration with the same
    transformation; the transformation aligns the views."""
    K = intrinsic_of(small_kinect.depth_intr)
    poses = syn.orbit_trajectory(8, yaw_sweep_deg=16.0)
    Rg = o3d.pipelines.registration
    voxel = 0.06

    def view(p):
        d, c = small_kinect.render_depth_aligned_rgb(p)
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c), o3d.geometry.Image(d), 1000.0, 3.0, False)
        pc = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, K)
        down = pc.voxel_down_sample(voxel)
        down.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=voxel * 2, max_nn=30))
        return down, Rg.compute_fpfh_feature(down, o3d.geometry.KDTreeSearchParamHybrid(radius=voxel * 5, max_nn=100))

    (src, fs), (tgt, ft) = view(poses[5]), view(poses[0])
    truth = np.linalg.inv(poses[0]) @ poses[5]                               # source camera -> target camera
    sx, tx = np.asarray(src.points), np.asarray(tgt.points)
    assert len(sx) > 500 and len(tx) > 500
    # nearest features: exact
    want_ij = o3r.match_features(fs.tensor.cpu().numpy(), ft.tensor.cpu().numpy())
    ij = torch.empty((len(sx),), dtype=torch.int32, device="cuda")
    _lib.check(_lib.load().b2r_match_features(_lib.ptr(fs.tensor), len(sx), _lib.ptr(ft.tensor), len(tx), 33, _lib.ptr(ij), _lib.stream_ptr()))
    assert np.array_equal(ij.cpu().numpy(), want_ij)
    corres = np.stack([np.arange(len(sx), dtype=np.int32), want_ij], 1)
    dist = voxel * 1.5
    crit = Rg.RANSACConvergenceCriteria(30000, 0.999)
    checkers = [Rg.CorrespondenceCheckerBasedOnEdgeLength(0.9), Rg.CorrespondenceCheckerBasedOnDistance(dist)]
    res = Rg.registration_ransac_based_on_correspondence(src, tgt, o3d.utility.Vector2iVector(corres), dist * 2, Rg.TransformationEstimationPointToPoint(False),
                                                         3, checkers, crit, seed=3)
    T_r, fit_r, rmse_r, rep_r = o3r.ransac_correspondence(sx, tx, corres, dist * 2, ransac_n=3, edge_thr=0.9, dist_thr=dist, max_iteration=30000,
                                                          confidence=0.999, seed=3)
    rep = res.ransac_report
    assert (rep["iterations"], rep["validated"], rep["winning_iteration"], rep["inlier_correspondences"]) == rep_r
    assert np.abs(res.transformation - T_r).max() <= 1e-12                   # same draws, same closed-form fit
    assert res.fitness == fit_r and abs(res.inlier_rmse - rmse_r) <= 1e-12
    assert res.fitness > 0.5
    # a coarse alignment (FPFH on a box room is ambiguous; max_correspondence_distance is 3 voxels) ...
    assert np.abs(res.transformation[:3, 3] - truth[:3, 3]).max() < 0.3 and np.abs(res.transformation[:3, :3] - truth[:3, :3]).max() < 0.3
    # ... which the reference then hands to point-to-plane ICP (BetterVisualizer2CamWorksCOLOR.py:252-259)
    fine = Rg.registration_icp(src, tgt, dist, res.transformation, Rg.TransformationEstimationPointToPlane(), Rg.ICPConvergenceCriteria(max_iteration=50))
    assert fine.fitness > 0.5
    assert np.abs(fine.transformation[:3, 3] - truth[:3, 3]).max() <= np.abs(res.transformation[:3, 3] - truth[:3, 3]).max() + 1e-3
    cs = np.asarray(res.correspondence_set)
    assert len(cs) == round(res.fitness * len(sx)) and (cs[:, 0] < len(sx)).all() and (cs[:, 1] < len(tx)).all()
    # the reference's call, verbatim argument style (older signature: no mutual_filter, "500" as the second criterion)
    ref_style = Rg.registration_ransac_based_on_feature_matching(
        src, tgt, fs, ft, max_correspondence_distance=dist * 2, estimation_method=Rg.TransformationEstimationPointToPoint(False), ransac_n=4,
        checkers=[Rg.CorrespondenceCheckerBasedOnEdgeLength(0.9), Rg.CorrespondenceCheckerBasedOnDistance(dist)],
        criteria=Rg.RANSACConvergenceCriteria(200000, 500))
    assert ref_style.ransac_report["iterations"] == 200000                   # confidence clamps to 1: the whole budget runs
    assert ref_style.fitness > 0.5 and ref_style.transformation.trace() != 4.0
    new_style = Rg.registration_ransac_based_on_feature_matching(src, tgt, fs, ft, True, dist * 2, Rg.TransformationEstimationPointToPoint(False), 3,
                                                                 checkers, Rg.RANSACConvergenceCriteria(20000, 0.999))
    assert new_style.fitness > 0.3
    # the validation primitive on its own == brute force
    from azurekinect3dreconstruction_b200.geometry import _Knn
    Ts = np.stack([truth, np.eye(4), res.transformation])
    knn = _Knn(dev(tx), dist * 2, 1)
    cnt = torch.empty((3,), dtype=torch.int64, device="cuda"); sm = torch.empty((3,), dtype=torch.float64, device="cuda")
    _lib.check(_lib.load().b2r_knn_evaluate_transforms(knn._h, _lib.ptr(dev(sx)), len(sx), _lib.ptr(dev(Ts)), 3, dist * 2, _lib.ptr(cnt), _lib.ptr(sm), _lib.stream_ptr()))
    torch.cuda.synchronize(); knn.close()
    for h in range(3):
        moved = o3r.transform_points(sx, Ts[h])
        d2 = ((moved[:, None, :] - tx[None]) ** 2).sum(-1).min(1)
        assert int(cnt[h]) == int((d2 < (dist * 2) ** 2).sum())
        assert abs(float(sm[h]) - d2[d2 < (dist * 2) ** 2].sum()) <= 1e-9
    with pytest.raises(RuntimeError):
        Rg.registration_ransac_based_on_correspondence(src, tgt, corres, dist, Rg.TransformationEstimationPointToPlane(), 3, [], crit)
