"""Committed golden fixtures (tests/golden/hotpath_v1.npz, made by tools/make_golden.py from the CPU oracle).

* not-gpu: the oracle still reproduces them (pins the oracle against accidental change);
* gpu: the CUDA path, through the C-ABI shim, reproduces them from the stored INPUTS (no renderer involved).
The reference itself holds no golden vectors and Open3D is not importable: parity vs real
Open3D stays unpinned.
"""
import os

import numpy as np
import pytest
import torch

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "hotpath_v1.npz"))
W, H = int(G["intr"][0]), int(G["intr"][1])
K = tuple(float(v) for v in G["intr"][2:6])
VOXEL, TRUNC = float(G["voxel"][0]), float(G["voxel"][1])


def canon_mesh(vtx, tri):
    o = np.lexsort(vtx.T[::-1])
    inv = np.empty_like(o)
    inv[o] = np.arange(len(o))
    t = inv[tri]
    return vtx[o], t[np.lexsort(t.T[::-1])].astype(np.int32)


# ---------------------------------------------------------------------------------------------- oracle vs golden
def test_oracle_reproduces_golden_unprojection(o3r):
    d0 = o3r.depth_to_float(G["depth"][0])
    assert np.array_equal(d0, G["a1_depth_f32"])
    assert np.array_equal(o3r.rgb_to_intensity(G["rgb"][0]), G["a1_intensity"])
    xyz, col, idx = o3r.unproject(d0, K, extrinsic=np.linalg.inv(G["poses"][0]), rgb=G["rgb"][0])
    assert np.array_equal(idx, G["a3_index"]) and np.array_equal(xyz, G["a3_xyz"]) and np.array_equal(col, G["a3_rgb"])


def oracle_volume(o3r):
    vol = o3r.TSDF(VOXEL, TRUNC, 1, parallel_units=True)
    for i in range(3):
        vol.integrate(o3r.depth_to_float(G["depth"][i]), K, np.linalg.inv(G["poses"][i]), rgb=G["rgb"][i])
    return vol


def test_oracle_reproduces_golden_tsdf_and_extraction(o3r):
    vol = oracle_volume(o3r)
    keys = vol.keys()
    keys = keys[np.lexsort(keys.T[::-1])]
    assert np.array_equal(keys, G["a8_keys"])
    for n, k in enumerate(keys):
        t, w, _ = vol.unit(k)
        assert float(t.astype(np.float64).sum()) == G["a9_tsdf_sum"][n] and float(w.astype(np.float64).sum()) == G["a9_weight_sum"][n]
    for j, n in enumerate(G["a9_sel"]):
        t, w, c = vol.unit(keys[n])
        assert np.array_equal(t, G["a9_tsdf"][j]) and np.array_equal(w, G["a9_weight"][j])
        assert np.array_equal(c.astype(np.float32), G["a9_rgb"][j])
    pts, _, nrm, _ = vol.extract_points()
    o = np.lexsort(pts.T[::-1])
    assert np.array_equal(pts[o], G["a10_xyz"]) and np.allclose(nrm[o], G["a10_nrm"], atol=1e-12)
    vtx, _, tri, _ = vol.extract_mesh()
    v, t = canon_mesh(vtx, tri)
    assert np.array_equal(v, G["a11_vtx"]) and np.array_equal(t, G["a11_tri"])


def test_oracle_reproduces_golden_odometry_and_icp(o3r):
    g = [o3r.rgb_to_intensity(G["rgb"][i]) for i in range(2)]
    d = [o3r.depth_to_float(G["depth"][i]) for i in range(2)]
    for jac in (0, 1):
        ok, T, info, _ = o3r.odometry(g[1], d[1], g[0], d[0], K, jac_kind=jac)
        assert int(ok) == int(G["a12_ok_%d" % jac][0])
        assert np.abs(T - G["a12_T_%d" % jac]).max() <= 1e-9           # OpenMP reduction order may differ
        assert np.abs(info - G["a12_info_%d" % jac]).max() <= 1e-6 * np.abs(info).max()
    T, fit, rmse, corr, it = o3r.icp_point_to_plane(G["a13_src"], G["a13_tgt"], G["a13_nrm"], 0.05, G["a13_init"], max_iter=30)
    assert it == int(G["a13_iters"][0]) and np.abs(T - G["a13_T"]).max() <= 1e-9
    assert np.array_equal(corr, G["a13_corr"]) and abs(fit - G["a13_fit"][0]) <= 1e-12


# ---------------------------------------------------------------------------------------------- CUDA vs golden
@pytest.mark.gpu
def test_cuda_reproduces_golden():
    import azurekinect3dreconstruction_b200 as o3d
    Ko = o3d.camera.PinholeCameraIntrinsic(W, H, *K)
    mk = lambda i, intensity=False: o3d.geometry.RGBDImage.create_from_color_and_depth(
        o3d.geometry.Image(G["rgb"][i]), o3d.geometry.Image(G["depth"][i]), 1000.0, 3.0, intensity)
    # a1 / a3
    rgbd = mk(0)
    assert np.array_equal(np.asarray(rgbd.depth), G["a1_depth_f32"])
    pcd = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, Ko, np.linalg.inv(G["poses"][0]))
    assert np.array_equal(pcd.source_index.cpu().numpy(), G["a3_index"])
    assert np.array_equal(np.asarray(pcd.points), G["a3_xyz"]) and np.array_equal(np.asarray(pcd.colors), G["a3_rgb"])
    # a8 / a9
    vol = o3d.pipelines.integration.ScalableTSDFVolume(VOXEL, TRUNC, o3d.pipelines.integration.TSDFVolumeColorType.RGB8, capacity_units=2048)
    for i in range(3):
        vol.integrate(mk(i), Ko, np.linalg.inv(G["poses"][i]))
    keys, tsdf, wgt, rgb = vol.export_units()
    keys = keys.cpu().numpy()
    o = np.lexsort(keys.T[::-1])
    assert np.array_equal(keys[o], G["a8_keys"])                       # block allocation: bit-exact
    tsdf, wgt, rgb = tsdf.cpu().numpy()[o], wgt.cpu().numpy()[o], rgb.cpu().numpy()[o]
    assert np.array_equal(tsdf.astype(np.float64).sum(1), G["a9_tsdf_sum"]) and np.array_equal(wgt.astype(np.float64).sum(1), G["a9_weight_sum"])
    for j, n in enumerate(G["a9_sel"]):
        assert np.array_equal(tsdf[n], G["a9_tsdf"][j]) and np.array_equal(wgt[n], G["a9_weight"][j])
        assert np.abs(rgb[n].T - G["a9_rgb"][j]).max() < 1e-3          # f32 running mean vs the oracle's f64
    # a10 / a11
    pc = vol.extract_point_cloud()
    pts = np.asarray(pc.points)
    o = np.lexsort(pts.T[::-1])
    assert pts.shape == G["a10_xyz"].shape and np.abs(pts[o] - G["a10_xyz"]).max() <= 1e-4 * VOXEL
    mesh = vol.extract_triangle_mesh()
    v, t = canon_mesh(np.asarray(mesh.vertices), np.asarray(mesh.triangles))
    assert np.array_equal(t, G["a11_tri"]) and np.abs(v - G["a11_vtx"]).max() <= 1e-4 * VOXEL
    # a12
    for jac, J in ((0, o3d.pipelines.odometry.RGBDOdometryJacobianFromColorTerm()), (1, o3d.pipelines.odometry.RGBDOdometryJacobianFromHybridTerm())):
        ok, T, info = o3d.pipelines.odometry.compute_rgbd_odometry(mk(1, True), mk(0, True), Ko, np.eye(4), J, o3d.pipelines.odometry.OdometryOption())
        assert int(ok) == int(G["a12_ok_%d" % jac][0])
        assert np.abs(T[:3, 3] - G["a12_T_%d" % jac][:3, 3]).max() <= 1e-6
        R = T[:3, :3] @ G["a12_T_%d" % jac][:3, :3].T
        assert np.arccos(np.clip((np.trace(R) - 1) / 2, -1, 1)) <= 1e-6
    # a13
    s, tg = o3d.geometry.PointCloud(G["a13_src"]), o3d.geometry.PointCloud(G["a13_tgt"])
    tg.normals = o3d.utility.Vector3dVector(G["a13_nrm"])
    res = o3d.pipelines.registration.registration_icp(s, tg, 0.05, G["a13_init"], o3d.pipelines.registration.TransformationEstimationPointToPlane(),
                                                       o3d.pipelines.registration.ICPConvergenceCriteria(max_iteration=30))
    assert np.abs(res.transformation - G["a13_T"]).max() <= 1e-6 and abs(res.fitness - G["a13_fit"][0]) <= 1e-9


# ---------------------------------------------------------------------------------------------- v2: a6 + f1
G2 = np.load(os.path.join(os.path.dirname(__file__), "golden", "hotpath_v2.npz"))


def test_oracle_reproduces_golden_v2_gather_and_pointcloud(o3r):
    Kc = tuple(float(v) for v in G2["a6_Kc"][2:6])
    gi = o3r.color_gather_index(G2["a6_depth"], K, int(G2["a6_Kc"][0]), int(G2["a6_Kc"][1]), Kc, G2["a6_d2c"])
    assert np.array_equal(gi.astype(np.int32), G2["a6_gather_index"])          # the a6 chain is defined by the oracle: pinned here
    xyz = G2["f1_xyz"]
    idx, d2, _ = o3r.pcd_search(xyz, None, 12, -1.0)
    assert np.array_equal(idx, G2["f1_knn_idx"]) and np.array_equal(d2, G2["f1_knn_d2"])
    hidx, _, hcnt = o3r.pcd_search(xyz, None, 16, 0.08)
    assert np.array_equal(hidx, G2["f1_hybrid_idx"]) and np.array_equal(hcnt, G2["f1_hybrid_cnt"])
    nrm, cov = o3r.pcd_estimate_normals(xyz, 16, 0.08)
    assert np.array_equal(cov, G2["f1_cov"]) and np.abs(nrm - G2["f1_normals"]).max() <= 1e-12
    keep, avg = o3r.pcd_statistical_outlier(xyz, 12, 2.0)
    assert np.array_equal(keep, G2["f1_keep"]) and np.array_equal(avg, G2["f1_avg"])
    vp, vc, _, vk = o3r.pcd_voxel_down_sample(xyz, 0.05, G2["f1_rgb"])
    assert np.array_equal(vk, G2["f1_voxel_key"]) and np.array_equal(vp, G2["f1_voxel_xyz"]) and np.array_equal(vc, G2["f1_voxel_rgb"])


@pytest.mark.gpu
def test_cuda_reproduces_golden_v2():
    import azurekinect3dreconstruction_b200 as o3d
    Kd = o3d.camera.PinholeCameraIntrinsic(W, H, *K)
    Kc = o3d.camera.PinholeCameraIntrinsic(int(G2["a6_Kc"][0]), int(G2["a6_Kc"][1]), *[float(v) for v in G2["a6_Kc"][2:6]])
    for dt in (torch.float64, torch.float32):                                   # generic (f64) and single-pass (f32) kernels
        pcd = o3d.geometry.PointCloud.create_from_kinect_capture(G2["a6_depth"], G2["a6_color"], Kd, Kc, G2["a6_d2c"], dtype=dt)
        src = pcd.source_index.cpu().numpy()
        assert np.array_equal(pcd.gather_index.cpu().numpy(), G2["a6_gather_index"].reshape(-1)[src])
    xyz = G2["f1_xyz"]
    pcd = o3d.geometry.PointCloud(xyz)
    pcd.colors = o3d.utility.Vector3dVector(G2["f1_rgb"])
    down = pcd.voxel_down_sample(0.05)
    assert np.array_equal(down.voxel_index.cpu().numpy(), G2["f1_voxel_key"])
    assert np.array_equal(np.asarray(down.points), G2["f1_voxel_xyz"]) and np.array_equal(np.asarray(down.colors), G2["f1_voxel_rgb"])
    inl, ind = pcd.remove_statistical_outlier(12, 2.0)
    assert np.array_equal(ind.cpu().numpy(), np.nonzero(G2["f1_keep"])[0])
    pcd.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(0.08, 16))
    dots = np.einsum("ij,ij->i", np.asarray(pcd.normals), G2["f1_normals"])
    assert np.abs(np.abs(dots) - 1.0).max() <= 1e-9


G3 = np.load(os.path.join(os.path.dirname(__file__), "golden", "hotpath_v3.npz"))


def _g3_pose_graph():
    import azurekinect3dreconstruction_b200 as o3d
    R = o3d.pipelines.registration
    pg = R.PoseGraph()
    for p in G3["f3_nodes"]:
        pg.nodes.append(R.PoseGraphNode(p))
    for k in range(len(G3["f3_src"])):
        pg.edges.append(R.PoseGraphEdge(int(G3["f3_src"][k]), int(G3["f3_dst"][k]), G3["f3_T"][k], G3["f3_info"][k], bool(G3["f3_uncertain"][k])))
    o = G3["f3_option"]
    return pg, R, R.GlobalOptimizationOption(float(o[0]), float(o[1]), float(o[2]), int(o[3]))


def test_golden_v3_pose_graph_and_cpu_oracles(o3r):
    """Rows f3 / f4: the CPU restatements reproduce the committed fixture (pins them against drift), and the host LM of
    libb2recon.so (envelope Cholesky) lands on the fixture's dense-solver result."""
    pg, R, opt = _g3_pose_graph()
    R.global_optimization(pg, R.GlobalOptimizationLevenbergMarquardt(), R.GlobalOptimizationConvergenceCriteria(), opt)
    rep = pg.last_report
    assert [rep["iterations"][0], rep["lm_steps"][0], rep["iterations"][1], rep["lm_steps"][1]] == G3["f3_iters"].tolist()
    assert np.array_equal(np.array([[e.source_node_id, e.target_node_id] for e in pg.edges], np.int32), G3["f3_out_edges"])
    assert np.abs(np.stack([nd.pose for nd in pg.nodes]) - G3["f3_out_nodes"]).max() <= 1e-9
    assert np.abs(np.array([e.confidence for e in pg.edges]) - G3["f3_out_conf"]).max() <= 1e-9
    T, fit, rmse, corr, it, grad = o3r.icp_colored(G3["c_src"], G3["c_src_rgb"], G3["c_tgt"], G3["c_tgt_nrm"], G3["c_tgt_rgb"], 0.03, 0.968, np.eye(4), max_iter=30)
    assert it == int(G3["c_iters"][0]) and np.array_equal(corr, G3["c_corr"])
    assert np.abs(T - G3["c_T"]).max() <= 1e-12 and np.abs(grad - G3["c_grad"]).max() <= 1e-9
    assert np.abs(T[:3, 3] - G3["c_truth"][:3, 3]).max() < 1e-3                 # the fixture itself is near the true motion
    fa = o3r.compute_fpfh(G3["g_src"], G3["g_src_nrm"], 60, 0.12)
    assert np.array_equal(fa, G3["g_fpfh_src"])
    ij = o3r.match_features(G3["g_fpfh_src"], G3["g_fpfh_tgt"])
    assert np.array_equal(ij, G3["g_match"])
    corres = np.stack([np.arange(len(ij), dtype=np.int32), ij], 1)
    RT, rfit, rrmse, rrep = o3r.ransac_correspondence(G3["g_src"], G3["g_tgt"], corres, 0.05, ransac_n=3, edge_thr=0.9, dist_thr=0.03, max_iteration=50000,
                                                      confidence=0.999, seed=5)
    assert list(rrep) == G3["g_report"].tolist() and np.array_equal(RT, G3["g_T"]) and rfit == G3["g_fit"][0]
    assert np.abs(RT - G3["g_truth"]).max() < 0.05


@pytest.mark.gpu
def test_cuda_reproduces_golden_v3():
    import azurekinect3dreconstruction_b200 as o3d
    R = o3d.pipelines.registration
    s, t = o3d.geometry.PointCloud(G3["c_src"]), o3d.geometry.PointCloud(G3["c_tgt"])
    s.colors, t.colors = o3d.utility.Vector3dVector(G3["c_src_rgb"]), o3d.utility.Vector3dVector(G3["c_tgt_rgb"])
    t.normals = o3d.utility.Vector3dVector(G3["c_tgt_nrm"])
    res = R.registration_colored_icp(s, t, 0.03, np.eye(4), R.TransformationEstimationForColoredICP(0.968), R.ICPConvergenceCriteria(1e-6, 1e-6, 30))
    assert res.num_iterations == int(G3["c_iters"][0])
    assert np.abs(res.transformation - G3["c_T"]).max() <= 1e-6 and np.abs(res.color_gradient.cpu().numpy() - G3["c_grad"]).max() <= 1e-9
    corr = np.full(len(G3["c_src"]), -1, np.int32)
    cs = np.asarray(res.correspondence_set)
    corr[cs[:, 0]] = cs[:, 1]
    assert np.array_equal(corr, G3["c_corr"])
    a, b = o3d.geometry.PointCloud(G3["g_src"]), o3d.geometry.PointCloud(G3["g_tgt"])
    a.normals, b.normals = o3d.utility.Vector3dVector(G3["g_src_nrm"]), o3d.utility.Vector3dVector(G3["g_tgt_nrm"])
    par = o3d.geometry.KDTreeSearchParamHybrid(radius=0.12, max_nn=60)
    fa, fb = R.compute_fpfh_feature(a, par), R.compute_fpfh_feature(b, par)
    ga, gb = fa.tensor.cpu().numpy(), fb.tensor.cpu().numpy()
    assert (np.abs(ga - G3["g_fpfh_src"]).max(1) > 1e-9).sum() <= 1 and (np.abs(gb - G3["g_fpfh_tgt"]).max(1) > 1e-9).sum() <= 1
    fa.tensor = torch.as_tensor(G3["g_fpfh_src"]).cuda()                       # the fixture's features: the rest must then be exact
    fb.tensor = torch.as_tensor(G3["g_fpfh_tgt"]).cuda()
    out = R.registration_ransac_based_on_feature_matching(a, b, fa, fb, False, 0.05, R.TransformationEstimationPointToPoint(False), 3,
                                                          [R.CorrespondenceCheckerBasedOnEdgeLength(0.9), R.CorrespondenceCheckerBasedOnDistance(0.03)],
                                                          R.RANSACConvergenceCriteria(50000, 0.999), seed=5)
    rep = out.ransac_report
    assert [rep["iterations"], rep["validated"], rep["winning_iteration"], rep["inlier_correspondences"]] == G3["g_report"].tolist()
    assert np.abs(out.transformation - G3["g_T"]).max() <= 1e-12 and out.fitness == G3["g_fit"][0]
