#!/usr/bin/env python3
"""Secondary measurements for the SURVEY 8(a) rows that are not the headline metric (bench.py is the contract):
unprojection (configs[0]/[1]), extraction (configs[2] tail), odometry + ICP (configs[3]).  CUDA-event timing,
warm-up first, inputs larger than L2 where the op streams.  Prints one JSON object.

    python tools/bench_ops.py [--frames 64] [--pairs 20] [--cpu]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import azurekinect3dreconstruction_b200 as o3d  # noqa: E402
from azurekinect3dreconstruction_b200 import geometry, synthetic as syn  # noqa: E402


def ev_time(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--pairs", type=int, default=20)
    ap.add_argument("--only-unproject", action="store_true")
    ap.add_argument("--cpu", action="store_true", help="also time the CPU oracle on bounded samples")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    out = {"hbm_peak_gbs": peak}
    kin = syn.SyntheticKinect(seed=0, dropout=0.02, noise_sigma_mm=1.0, device="cuda")
    poses = syn.orbit_trajectory(300)
    di, ci = kin.depth_intr, kin.color_intr
    Kd = o3d.camera.PinholeCameraIntrinsic(di["width"], di["height"], di["fx"], di["fy"], di["cx"], di["cy"])
    Kc = o3d.camera.PinholeCameraIntrinsic(ci["width"], ci["height"], ci["fx"], ci["fy"], ci["cx"], ci["cy"])
    H, W = di["height"], di["width"]

    # ---- a1+a3..a6: unprojection, batch of B Kinect captures (native depth + 720p BGRA, colour gathered to depth)
    nsrc = 8
    caps = [kin.render_pose(poses[i * 30], want_transformed_depth=False) for i in range(nsrc)]
    B = a.batch
    depth = torch.stack([torch.from_numpy(np.ascontiguousarray(caps[i % nsrc].depth).view(np.int16)).to(dev) for i in range(B)]).contiguous().view(torch.uint16)
    color = torch.stack([torch.as_tensor(caps[i % nsrc].color).to(dev) for i in range(B)]).contiguous()
    d2c = kin.d2c
    flip = np.diag([1.0, -1.0, -1.0, 1.0])
    T = np.stack([flip @ poses[(i * 30) % 300] for i in range(B)])

    def unp_gather(b=B):
        return geometry.unproject(depth[:b], Kd, depth_is_u16=True, depth_scale=1000.0, depth_trunc=3.0, color=color[:b], color_bgr=True,
                                  color_intrinsic=Kc, depth_to_color=d2c, transforms=T[:b], dtype=torch.float32)
    r = unp_gather()
    n_valid = float(r["counts"].sum().item()) / B
    ms = ev_time(unp_gather, iters=5)
    algo = 2.0 * W * H + 4.0 * n_valid + 24.0 * n_valid
    out["unproject_kinect_gather_f32"] = {"batch": B, "ms_per_launch_sequence": ms, "frames_per_s": B / (ms * 1e-3), "valid_points_per_frame": n_valid,
                                          "algorithmic_bytes_per_frame": algo, "achieved_gbs": algo * B / (ms * 1e-3) / 1e9,
                                          "frac_of_measured_hbm": algo * B / (ms * 1e-3) / 1e9 / peak}
    # the same launch with the output buffers and the device copy of the transforms reused (what a streaming caller does):
    # one memset of the look-back words + ONE kernel per call -- this is the figure the roofline fraction is quoted on
    def unp_reuse():
        return geometry.unproject(depth, Kd, depth_is_u16=True, depth_scale=1000.0, depth_trunc=3.0, color=color, color_bgr=True,
                                  color_intrinsic=Kc, depth_to_color=d2c, dtype=torch.float32, out=r)
    ms_k = ev_time(unp_reuse, iters=10)
    out["unproject_kinect_gather_f32"].update({"ms_buffers_reused": ms_k, "frames_per_s_buffers_reused": B / (ms_k * 1e-3),
                                                "achieved_gbs_buffers_reused": algo * B / (ms_k * 1e-3) / 1e9,
                                                "frac_of_measured_hbm_buffers_reused": algo * B / (ms_k * 1e-3) / 1e9 / peak})
    ms1 = ev_time(lambda: unp_gather(1), iters=50, warm=5)
    ms2 = ev_time(lambda: unp_gather(2), iters=50, warm=5)
    out["unproject_kinect_gather_f32"]["latency_ms_B1"] = ms1
    out["two_camera_fusion_config2"] = {"ms_per_tick_B2": ms2, "ticks_per_s": 1e3 / ms2, "sensor_rate_hz": 30}
    # aligned colour, f64 output (bit-exact upstream chain)
    dal, cal = [], []
    for i in range(nsrc):
        d, c = kin.render_depth_aligned_rgb(poses[i * 30])
        dal.append(d.view(torch.int16).to(dev)); cal.append(c.to(dev))
    Bd = min(B, 64)
    d64 = torch.stack([dal[i % nsrc] for i in range(Bd)]).contiguous().view(torch.uint16)
    c64 = torch.stack([cal[i % nsrc] for i in range(Bd)]).contiguous()

    def unp_f64():
        return geometry.unproject(d64, Kd, depth_is_u16=True, depth_scale=1000.0, depth_trunc=3.0,
                                  color=c64, dtype=torch.float64)
    ms = ev_time(unp_f64, iters=5)
    out["unproject_aligned_f64"] = {"batch": Bd, "ms": ms, "frames_per_s": Bd / (ms * 1e-3)}
    del depth, color, r
    torch.cuda.empty_cache()
    if a.only_unproject:
        print(json.dumps(out, indent=1))
        return

    # ---- a10/a11: extraction after integrating --frames frames at 4 mm / 2 cm
    TT = o3d.pipelines.integration.TSDFVolumeColorType
    vol = o3d.pipelines.integration.ScalableTSDFVolume(0.004, 0.02, TT.RGB8, capacity_units=40000)
    F = a.frames
    fr = [kin.render_depth_aligned_rgb(poses[i]) for i in range(F)]
    dd = torch.stack([f[0].view(torch.int16).to(dev) for f in fr]).contiguous().view(torch.uint16)
    cc = torch.stack([f[1].to(dev) for f in fr]).contiguous()
    E = np.stack([np.linalg.inv(poses[i]) for i in range(F)])
    vol.integrate_raw(dd, cc, Kd, E, 1000.0, 3.0)
    units = vol.num_units()
    mesh = vol.extract_triangle_mesh()
    V, Fc = len(mesh.vertices), len(mesh.triangles)
    ms_mesh = ev_time(lambda: vol.extract_triangle_mesh(), iters=3, warm=1)
    # the same after the volume changed (one more frame integrated before every extraction: sort + count pass run again);
    # the integrate call alone is timed and subtracted
    one = lambda: vol.integrate_raw(dd[:1], cc[:1], Kd, E[:1], 1000.0, 3.0)
    ms_int = ev_time(lambda: (one(), vol.flush()), iters=3, warm=1)
    ms_mesh_changed = ev_time(lambda: (one(), vol.extract_triangle_mesh()), iters=3, warm=1) - ms_int
    pc = vol.extract_point_cloud()
    Kp = len(pc.points)
    ms_pts = ev_time(lambda: vol.extract_point_cloud(), iters=3, warm=1)
    ms_nrm = ev_time(lambda: mesh.compute_vertex_normals(), iters=3, warm=1)
    nvox = units * 4096.0
    algo_mesh = 8.0 * nvox + 24.0 * V + 24.0 * V + 12.0 * Fc
    algo_pts = 8.0 * nvox + 24.0 * Kp + 36.0 * Kp
    out["extract_triangle_mesh"] = {"units": units, "vertices": V, "triangles": Fc, "ms": ms_mesh, "algorithmic_bytes": algo_mesh,
                                    "achieved_gbs": algo_mesh / (ms_mesh * 1e-3) / 1e9, "frac_of_measured_hbm": algo_mesh / (ms_mesh * 1e-3) / 1e9 / peak,
                                    "ms_after_volume_change": ms_mesh_changed,
                                    "note": "ms: volume unchanged since the last extraction (sorted order / vertex slots cached, fill pass only); "
                                            "ms_after_volume_change: sort + count + scans + fill (two-call protocol, f64 outputs)"}
    out["extract_point_cloud"] = {"points": Kp, "ms": ms_pts, "algorithmic_bytes": algo_pts, "achieved_gbs": algo_pts / (ms_pts * 1e-3) / 1e9}
    out["compute_vertex_normals"] = {"ms": ms_nrm}

    # ---- a12/a13: odometry + ICP over consecutive pairs
    mk = lambda i: o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(cc[i]), o3d.geometry.Image(dd[i]), 1000.0, 3.0, True)
    imgs = [mk(i) for i in range(min(F, a.pairs + 1))]
    opt = o3d.pipelines.odometry.OdometryOption(iteration_number_per_pyramid_level=o3d.utility.IntVector([20, 10, 5]), max_depth_diff=0.03, min_depth=0.0, max_depth=4.0)
    J = o3d.pipelines.odometry.RGBDOdometryJacobianFromHybridTerm()
    res = []

    def odo():
        res.clear()
        for i in range(len(imgs) - 1):
            res.append(o3d.pipelines.odometry.compute_rgbd_odometry(imgs[i + 1], imgs[i], Kd, np.eye(4), J, opt))
    t0 = time.perf_counter(); odo(); torch.cuda.synchronize(); t_first = time.perf_counter() - t0
    t0 = time.perf_counter(); odo(); torch.cuda.synchronize(); t_odo = time.perf_counter() - t0
    errs = []
    for i, (ok, T_, _) in enumerate(res):
        gt = np.linalg.inv(poses[i]) @ poses[i + 1]
        errs.append(float(np.linalg.norm(T_[:3, 3] - gt[:3, 3])))
    out["rgbd_odometry_hybrid"] = {"pairs": len(res), "ms_per_pair": 1e3 * t_odo / max(1, len(res)), "pairs_per_s": len(res) / t_odo, "all_ok": all(r[0] for r in res),
                                   "median_translation_error_m_vs_ground_truth": float(np.median(errs)), "first_call_s": t_first}
    from _helpers import image_normals
    p_src = o3d.geometry.PointCloud.create_from_depth_image(o3d.geometry.Image(dd[1]), Kd, depth_scale=1000.0, depth_trunc=3.0, stride=4)
    p_tgt = o3d.geometry.PointCloud.create_from_depth_image(o3d.geometry.Image(dd[0]), Kd, depth_scale=1000.0, depth_trunc=3.0, stride=4)
    d0 = (dd[0]).cpu().numpy().astype(np.float32) / 1000.0
    d0[d0 >= 3.0] = 0
    p_tgt.normals = o3d.utility.Vector3dVector(image_normals(d0, (di["fx"], di["fy"], di["cx"], di["cy"]), stride=4))
    crit = o3d.pipelines.registration.ICPConvergenceCriteria(max_iteration=100)
    est = o3d.pipelines.registration.TransformationEstimationPointToPlane()
    icp = lambda: o3d.pipelines.registration.registration_icp(p_src, p_tgt, 0.03, np.eye(4), est, crit)
    r_icp = icp()
    t0 = time.perf_counter(); r_icp = icp(); torch.cuda.synchronize(); t_icp = time.perf_counter() - t0
    out["icp_point_to_plane"] = {"source_points": len(p_src.points), "target_points": len(p_tgt.points), "iterations": int(r_icp.num_iterations), "ms": 1e3 * t_icp,
                                 "fitness": float(r_icp.fitness), "inlier_rmse": float(r_icp.inlier_rmse)}

    # ---- f4 (part): registration_colored_icp between consecutive frames (GoodColorReconstruction.py:246-254)
    def colored_cloud(i):
        r_ = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(cc[i]), o3d.geometry.Image(dd[i]), 1000.0, 3.0, False)
        return o3d.geometry.PointCloud.create_from_rgbd_image(r_, Kd).voxel_down_sample(0.02)
    c_src, c_tgt = colored_cloud(1), colored_cloud(0)
    c_tgt.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=0.04, max_nn=30))
    c_tgt.orient_normals_towards_camera_location(np.array([0.0, 0.0, 0.0]))
    cicp = lambda: o3d.pipelines.registration.registration_colored_icp(c_src, c_tgt, 0.03, np.eye(4), o3d.pipelines.registration.TransformationEstimationForColoredICP(),
                                                                       o3d.pipelines.registration.ICPConvergenceCriteria(relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=100))
    r_c = cicp(); torch.cuda.synchronize()
    t0 = time.perf_counter(); r_c = cicp(); torch.cuda.synchronize(); t_c = time.perf_counter() - t0
    gt01 = np.linalg.inv(poses[0]) @ poses[1]
    out["colored_icp_f4"] = {"source_points": len(c_src.points), "target_points": len(c_tgt.points), "iterations": int(r_c.num_iterations), "ms": 1e3 * t_c,
                             "fitness": float(r_c.fitness), "inlier_rmse": float(r_c.inlier_rmse),
                             "translation_error_m_vs_ground_truth": float(np.linalg.norm(r_c.transformation[:3, 3] - gt01[:3, 3]))}

    # ---- f4: preprocess_point_cloud + global registration of two views (BetterVisualizer2CamWorksCOLOR.py:197-243), full NFOV frames
    Rg4 = o3d.pipelines.registration
    vox = 0.05
    def prep(i):
        r_ = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(cc[i]), o3d.geometry.Image(dd[i]), 1000.0, 3.0, False)
        pc_, _ = o3d.geometry.PointCloud.create_from_rgbd_image(r_, Kd).remove_statistical_outlier(nb_neighbors=20, std_ratio=2.0)
        dn_ = pc_.voxel_down_sample(vox)
        dn_.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=vox * 2, max_nn=30))
        torch.cuda.synchronize(); t0_ = time.perf_counter()
        f_ = Rg4.compute_fpfh_feature(dn_, o3d.geometry.KDTreeSearchParamHybrid(radius=vox * 5, max_nn=100)); torch.cuda.synchronize()
        return dn_, f_, time.perf_counter() - t0_
    far = min(len(poses) - 1, F - 1)
    prep(0)
    (g_src, g_fs, t_f1), (g_tgt, g_ft, t_f2) = prep(far), prep(0)
    dthr = vox * 1.5
    def ransac(iters, conf):
        return Rg4.registration_ransac_based_on_feature_matching(
            g_src, g_tgt, g_fs, g_ft, max_correspondence_distance=dthr * 2, estimation_method=Rg4.TransformationEstimationPointToPoint(False), ransac_n=4,
            checkers=[Rg4.CorrespondenceCheckerBasedOnEdgeLength(0.9), Rg4.CorrespondenceCheckerBasedOnDistance(dthr)],
            criteria=Rg4.RANSACConvergenceCriteria(iters, conf))
    ransac(100000, 0.999); torch.cuda.synchronize()
    t0 = time.perf_counter(); r_g = ransac(4000000, 500); torch.cuda.synchronize(); t_g = time.perf_counter() - t0      # the reference's criteria
    t0 = time.perf_counter(); r_q = ransac(4000000, 0.999); torch.cuda.synchronize(); t_q = time.perf_counter() - t0
    gtf = np.linalg.inv(poses[0]) @ poses[far]
    out["global_registration_f4"] = {"voxel": vox, "source_points": len(g_src.points), "target_points": len(g_tgt.points), "fpfh_ms": 1e3 * 0.5 * (t_f1 + t_f2),
                                     "ransac_4M_iterations_ms": 1e3 * t_g, "ransac_4M_report": r_g.ransac_report, "fitness": float(r_g.fitness),
                                     "translation_error_m_vs_ground_truth": float(np.linalg.norm(r_g.transformation[:3, 3] - gtf[:3, 3])),
                                     "ransac_confidence_0.999_ms": 1e3 * t_q, "ransac_confidence_0.999_report": r_q.ransac_report,
                                     "views": [0, far]}

    # ---- f1: per-frame conditioning chain of the reference (GoodColorReconstruction.py:141-153) on one full NFOV frame
    rgbd0 = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(cc[0]), o3d.geometry.Image(dd[0]), 1000.0, 3.0, False)
    full = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd0, Kd)
    nfull = len(full.points)

    def t_wall(fn, iters=3):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(iters):
            r_ = fn()
        torch.cuda.synchronize()
        return 1e3 * (time.perf_counter() - t0) / iters, r_
    ms_out, (inl, _) = t_wall(lambda: full.remove_statistical_outlier(nb_neighbors=20, std_ratio=2.0))
    ms_vox, down = t_wall(lambda: inl.voxel_down_sample(0.01))
    ms_nrm, _ = t_wall(lambda: down.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=0.1, max_nn=30)))
    ms_nrm2, _ = t_wall(lambda: down.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=0.02, max_nn=30)))
    ms_ori, _ = t_wall(lambda: down.orient_normals_towards_camera_location(np.array([0.0, 0.0, 0.0])))
    out["pointcloud_conditioning_f1"] = {"input_points": nfull, "remove_statistical_outlier_20_2.0_ms": ms_out, "inliers": len(inl.points),
                                         "voxel_down_sample_1cm_ms": ms_vox, "voxels": len(down.points),
                                         "estimate_normals_hybrid_r0.1_nn30_ms": ms_nrm, "estimate_normals_hybrid_r0.02_nn30_ms": ms_nrm2,
                                         "orient_normals_ms": ms_ori, "timing": "wall clock around the public call incl. index build and result allocation"}

    # ---- f3: pose graph from the odometry results above + a closure -> global_optimization (host LM) -> re-integration
    Rg = o3d.pipelines.registration
    pg = Rg.PoseGraph()
    pg.nodes.append(Rg.PoseGraphNode(poses[0]))
    cur = poses[0].copy()
    for i, (ok, T_, info_) in enumerate(res):
        cur = cur @ T_
        pg.nodes.append(Rg.PoseGraphNode(cur))
        pg.edges.append(Rg.PoseGraphEdge(i, i + 1, np.linalg.inv(T_), info_, uncertain=False))
    nn = len(pg.nodes)
    pg.edges.append(Rg.PoseGraphEdge(0, nn - 1, np.linalg.inv(poses[nn - 1]) @ poses[0], res[0][2], uncertain=True))
    t0 = time.perf_counter()
    Rg.global_optimization(pg, Rg.GlobalOptimizationLevenbergMarquardt(), Rg.GlobalOptimizationConvergenceCriteria(),
                           Rg.GlobalOptimizationOption(max_correspondence_distance=0.02, edge_prune_threshold=0.25, preference_loop_closure=1.0, reference_node=0))
    t_go = time.perf_counter() - t0
    vol3 = o3d.pipelines.integration.ScalableTSDFVolume(0.004, 0.02, o3d.pipelines.integration.TSDFVolumeColorType.RGB8)
    o3d.pipelines.integration.reintegrate_frames(vol3, dd[:nn], cc[:nn], pg, Kd); torch.cuda.synchronize()
    t0 = time.perf_counter()
    o3d.pipelines.integration.reintegrate_frames(vol3, dd[:nn], cc[:nn], pg, Kd); vol3.flush(); torch.cuda.synchronize()
    t_re = time.perf_counter() - t0
    # a long synthetic chain for the solver alone: 2000 nodes, closures every 100
    rngp = np.random.default_rng(0)
    big = Rg.PoseGraph(); traj = syn.orbit_trajectory(2000); cur = traj[0].copy(); big.nodes.append(Rg.PoseGraphNode(cur))
    I6 = np.identity(6) * 2000.0
    for i in range(1, 2000):
        v = np.concatenate([rngp.normal(0, 5e-4, 3), rngp.normal(0, 5e-4, 3)])
        dT = np.identity(4); dT[:3, 3] = v[3:]; dT[0, 1], dT[1, 0], dT[0, 2], dT[2, 0], dT[1, 2], dT[2, 1] = -v[2], v[2], v[1], -v[1], -v[0], v[0]
        T_ = np.linalg.inv(traj[i]) @ traj[i - 1] @ dT
        cur = cur @ np.linalg.inv(T_); big.nodes.append(Rg.PoseGraphNode(cur)); big.edges.append(Rg.PoseGraphEdge(i - 1, i, T_, I6, False))
    for i in range(0, 1800, 100):
        big.edges.append(Rg.PoseGraphEdge(i, i + 150, np.linalg.inv(traj[i + 150]) @ traj[i], I6, True))
    t0 = time.perf_counter()
    Rg.global_optimization(big, Rg.GlobalOptimizationLevenbergMarquardt(), Rg.GlobalOptimizationConvergenceCriteria(),
                           Rg.GlobalOptimizationOption(max_correspondence_distance=0.05, edge_prune_threshold=0.25, preference_loop_closure=2.0, reference_node=0))
    t_big = time.perf_counter() - t0
    out["posegraph_f3"] = {"nodes": nn, "edges": len(pg.edges), "global_optimization_ms": 1e3 * t_go, "report": pg.last_report,
                           "reintegrate_frames": nn, "reintegrate_ms": 1e3 * t_re, "reintegrate_frames_per_s": nn / t_re,
                           "chain_2000_nodes_18_closures_ms": 1e3 * t_big, "chain_2000_report": big.last_report,
                           "note": "solver = host block-envelope LM inside libb2recon.so; re-integration = reset + batched K1/K2 on the GPU"}

    if a.cpu:   # bounded CPU-oracle samples (reported baseline only)
        from oracle import o3r
        Kt = (di["fx"], di["fy"], di["cx"], di["cy"])
        dn = (dd[0]).cpu().numpy()
        cn = cc[0].cpu().numpy()
        t0 = time.perf_counter(); o3r.unproject(o3r.depth_to_float(dn), Kt, rgb=cn); t_u = time.perf_counter() - t0
        g = [o3r.rgb_to_intensity(cc[i].cpu().numpy()) for i in range(2)]
        df = [o3r.depth_to_float((dd[i]).cpu().numpy()) for i in range(2)]
        t0 = time.perf_counter(); o3r.odometry(g[1], df[1], g[0], df[0], Kt, jac_kind=1); t_o = time.perf_counter() - t0
        cv = o3r.TSDF(0.004, 0.02, 1, parallel_units=True)
        nf = min(F, 8)
        for i in range(nf):
            cv.integrate(o3r.depth_to_float((dd[i]).cpu().numpy()), Kt, E[i], rgb=cc[i].cpu().numpy())
        t0 = time.perf_counter(); vtx, _, tri, _ = cv.extract_mesh(); t_m = time.perf_counter() - t0
        out["cpu_oracle"] = {"cores": o3r.num_threads(), "unproject_one_frame_ms": 1e3 * t_u, "odometry_one_pair_ms": 1e3 * t_o,
                             "extract_mesh_ms": 1e3 * t_m, "extract_mesh_units": cv.num_units(), "extract_mesh_vertices": int(len(vtx)), "kind": "port"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
