#!/usr/bin/env python3
"""Generates tests/golden/hotpath_v1.npz: small seeded inputs + the CPU oracle's outputs for every row of the hot
path.  The reference holds no golden vectors of its own and Open3D cannot be
imported here, so these fixtures pin the ORACLE against accidental change and give the GPU tests inputs that
do not depend on the synthetic renderer's floating-point behaviour on the machine that runs them.

    python tools/make_golden.py          # rewrites tests/golden/hotpath_v1.npz
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from azurekinect3dreconstruction_b200 import synthetic as syn  # noqa: E402
from oracle import o3r  # noqa: E402
from _helpers import image_normals  # noqa: E402

INTR = dict(width=80, height=72, fx=504.2 / 8, fy=504.1 / 8, cx=319.8 / 8, cy=287.9 / 8)
VOXEL, TRUNC = 0.02, 0.08


def main():
    if "--only-v3" in sys.argv:
        make_v3()
        return
    kin = syn.SyntheticKinect(depth_intrinsics=INTR, seed=7, dropout=0.02)
    poses = syn.orbit_trajectory(3, yaw_sweep_deg=4.0)
    K = (INTR["fx"], INTR["fy"], INTR["cx"], INTR["cy"])
    depth, rgb = [], []
    for p in poses:
        d, c = kin.render_depth_aligned_rgb(p)
        depth.append(d.view(torch.int16).numpy().view(np.uint16).copy())
        rgb.append(c.numpy().copy())
    depth, rgb, poses = np.stack(depth), np.stack(rgb), np.stack(poses)
    out = dict(intr=np.array([INTR["width"], INTR["height"], *K], np.float64), voxel=np.array([VOXEL, TRUNC]),
               depth=depth, rgb=rgb, poses=poses)
    # a1 + a3
    d0 = o3r.depth_to_float(depth[0])
    E0 = np.linalg.inv(poses[0])
    xyz, col, idx = o3r.unproject(d0, K, extrinsic=E0, rgb=rgb[0])
    out.update(a1_depth_f32=d0, a1_intensity=o3r.rgb_to_intensity(rgb[0]), a3_xyz=xyz, a3_rgb=col, a3_index=idx)
    # a8 + a9
    vol = o3r.TSDF(VOXEL, TRUNC, 1, parallel_units=True)
    for i in range(3):
        vol.integrate(o3r.depth_to_float(depth[i]), K, np.linalg.inv(poses[i]), rgb=rgb[i])
    keys = vol.keys()
    order = np.lexsort(keys.T[::-1])
    keys = keys[order]
    tsdf_sum, w_sum, planes_t, planes_w, planes_c = [], [], [], [], []
    for n, k in enumerate(keys):
        t, w, c = vol.unit(k)
        tsdf_sum.append(float(t.astype(np.float64).sum()))
        w_sum.append(float(w.astype(np.float64).sum()))
        if n % max(1, len(keys) // 6) == 0 and len(planes_t) < 6:
            planes_t.append(t); planes_w.append(w); planes_c.append(c.astype(np.float32))
    sel = [n for n in range(len(keys)) if n % max(1, len(keys) // 6) == 0][:6]
    out.update(a8_keys=keys, a9_tsdf_sum=np.array(tsdf_sum), a9_weight_sum=np.array(w_sum), a9_sel=np.array(sel, np.int32),
               a9_tsdf=np.stack(planes_t), a9_weight=np.stack(planes_w), a9_rgb=np.stack(planes_c))
    # a10 + a11 (canonical order: sort by xyz)
    pts, pcol, pnrm, _ = vol.extract_points()
    o = np.lexsort(pts.T[::-1])
    out.update(a10_xyz=pts[o], a10_nrm=pnrm[o])
    vtx, vcol, tri, _ = vol.extract_mesh()
    o = np.lexsort(vtx.T[::-1])
    inv = np.empty_like(o); inv[o] = np.arange(len(o))
    t2 = inv[tri]
    t2 = t2[np.lexsort(t2.T[::-1])]
    out.update(a11_vtx=vtx[o], a11_tri=t2.astype(np.int32))
    # a12
    g = [o3r.rgb_to_intensity(rgb[i]) for i in range(2)]
    df = [o3r.depth_to_float(depth[i]) for i in range(2)]
    for jac in (0, 1):
        ok, T, info, _ = o3r.odometry(g[1], df[1], g[0], df[0], K, jac_kind=jac)
        out["a12_T_%d" % jac] = T
        out["a12_info_%d" % jac] = info
        out["a12_ok_%d" % jac] = np.array([int(ok)])
    # a13
    src, _, _ = o3r.unproject(df[2] if len(df) > 2 else o3r.depth_to_float(depth[2]), K)
    tgt, _, _ = o3r.unproject(df[0], K)
    nrm = image_normals(df[0], K)
    init = np.linalg.inv(poses[0]) @ poses[2]
    init[:3, 3] += [0.003, -0.002, 0.002]
    T, fit, rmse, corr, it = o3r.icp_point_to_plane(src, tgt, nrm, 0.05, init, max_iter=30)
    out.update(a13_src=src, a13_tgt=tgt, a13_nrm=nrm, a13_init=init, a13_T=T, a13_fit=np.array([fit, rmse]), a13_iters=np.array([it]),
               a13_corr=corr)
    path = os.path.join(ROOT, "tests", "golden", "hotpath_v1.npz")
    if "--only-v2" not in sys.argv:
        np.savez_compressed(path, **out)
        print(path, os.path.getsize(path), "bytes;", len(keys), "units,", len(vtx), "vertices,", len(tri), "triangles")
    make_v2(kin, poses, depth, rgb, K)


def make_v2(kin, poses, depth, rgb, K):
    """tests/golden/hotpath_v2.npz: the rows added after v1 -- a6 (colour-to-depth gather on a native Kinect capture) and
    f1 (neighbour search, normals, outlier removal, voxel down-sampling) on the unprojected first frame."""
    cdict = dict(width=160, height=90, fx=605.6 / 8, fy=605.9 / 8, cx=637.7 / 8, cy=364.3 / 8)
    kin2 = syn.SyntheticKinect(depth_intrinsics=INTR, color_intrinsics=cdict, seed=7, dropout=0.02)
    cap = kin2.render_pose(poses[1], want_transformed_depth=False)
    Kc = (cdict["fx"], cdict["fy"], cdict["cx"], cdict["cy"])
    gi = o3r.color_gather_index(cap.depth, K, cdict["width"], cdict["height"], Kc, kin2.d2c)
    out = dict(a6_depth=np.ascontiguousarray(cap.depth), a6_color=np.ascontiguousarray(cap.color), a6_Kc=np.array([cdict["width"], cdict["height"], *Kc]),
               a6_d2c=kin2.d2c, a6_gather_index=gi.astype(np.int32))
    xyz, col, _ = o3r.unproject(o3r.depth_to_float(depth[0]), K, rgb=rgb[0])
    idx, d2, cnt = o3r.pcd_search(xyz, None, 12, -1.0)
    hidx, _, hcnt = o3r.pcd_search(xyz, None, 16, 0.08)
    nrm, cov = o3r.pcd_estimate_normals(xyz, 16, 0.08)
    keep, avg = o3r.pcd_statistical_outlier(xyz, 12, 2.0)
    vp, vc, _, vk = o3r.pcd_voxel_down_sample(xyz, 0.05, col)
    out.update(f1_xyz=xyz, f1_rgb=col, f1_knn_idx=idx, f1_knn_d2=d2, f1_hybrid_idx=hidx, f1_hybrid_cnt=hcnt, f1_cov=cov, f1_normals=nrm,
               f1_keep=keep, f1_avg=avg, f1_voxel_xyz=vp, f1_voxel_rgb=vc, f1_voxel_key=vk)
    path = os.path.join(ROOT, "tests", "golden", "hotpath_v2.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", len(xyz), "points,", int((gi >= 0).sum()), "gathered texels,", len(vp), "voxels")


def make_v3():
    """tests/golden/hotpath_v3.npz: rows f3 / f4 -- pose-graph optimisation, coloured ICP (gradients + pose), FPFH, nearest
    features and RANSAC, on inputs small enough for the dense / brute-force CPU restatements."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import posegraph as pgref
    rng = np.random.default_rng(11)
    # f3: 14-node orbit, noisy odometry chain, two true closures, one false closure
    n = 14
    gt = syn.orbit_trajectory(n)
    info = np.identity(6) * 2000.0
    mot = lambda r, t: pgref.transform_vector6d_to_matrix4d(np.concatenate([rng.normal(0, r, 3), rng.normal(0, t, 3)]))
    nodes, edges = [gt[0].copy()], []
    for i in range(1, n):
        T = np.linalg.inv(gt[i]) @ gt[i - 1] @ mot(0.002, 0.003)
        nodes.append(nodes[i - 1] @ np.linalg.inv(T))
        edges.append(dict(s=i - 1, t=i, T=T, info=info, uncertain=False, conf=1.0))
    for a, b in ((0, 13), (2, 10)):
        edges.append(dict(s=a, t=b, T=np.linalg.inv(gt[b]) @ gt[a], info=info, uncertain=True, conf=1.0))
    edges.append(dict(s=1, t=8, T=np.linalg.inv(gt[8]) @ gt[1] @ pgref.transform_vector6d_to_matrix4d(np.array([0.3, -0.2, 0.4, 0.5, -0.4, 0.3])),
                      info=info, uncertain=True, conf=1.0))
    opt = dict(max_correspondence_distance=0.05, edge_prune_threshold=0.25, preference_loop_closure=2.0, reference_node=0)
    on, oe, (r1, r2) = pgref.global_optimization(nodes, edges, opt)
    out = dict(f3_nodes=np.stack(nodes), f3_src=np.array([e["s"] for e in edges], np.int32), f3_dst=np.array([e["t"] for e in edges], np.int32),
               f3_T=np.stack([e["T"] for e in edges]), f3_info=np.stack([e["info"] for e in edges]),
               f3_uncertain=np.array([e["uncertain"] for e in edges], np.uint8), f3_option=np.array([0.05, 0.25, 2.0, 0.0]),
               f3_out_nodes=np.stack(on), f3_out_edges=np.array([[e["s"], e["t"]] for e in oe], np.int32),
               f3_out_conf=np.array([e["conf"] for e in oe]), f3_iters=np.array([r1[0], r1[1], r2[0], r2[1]], np.int64))
    # f4 coloured ICP: a gently curved textured sheet and a slightly moved, re-sampled copy
    def sheet(m, seed):
        r = np.random.default_rng(seed)
        u, v = r.random(m) * 0.5 - 0.25, r.random(m) * 0.5 - 0.25
        z = 2.0 + 0.03 * np.sin(5 * u) + 0.02 * v * v
        p = np.stack([u, v, z], 1)
        nz = np.stack([-0.15 * np.cos(5 * u), -0.04 * v, np.ones_like(u)], 1)
        inten = 0.5 + 0.25 * np.sin(18 * u) * np.cos(14 * v) + 0.1 * u
        return p, nz / np.linalg.norm(nz, axis=1, keepdims=True), np.stack([inten, inten * 0.9 + 0.05, inten * 0.8 + 0.1], 1)
    tp, tn, tc = sheet(1400, 1)
    sp, _, sc = sheet(900, 2)
    M = syn.make_pose(syn.rot_z(0.01) @ syn.rot_x(-0.008), (0.004, -0.003, 0.002))
    sp = o3r.transform_points(sp, np.linalg.inv(M))
    T, fit, rmse, corr, it, grad = o3r.icp_colored(sp, sc, tp, tn, tc, 0.03, 0.968, np.eye(4), max_iter=30)
    out.update(c_src=sp, c_src_rgb=sc, c_tgt=tp, c_tgt_nrm=tn, c_tgt_rgb=tc, c_T=T, c_fit=np.array([fit, rmse]), c_iters=np.array([it]),
               c_corr=corr, c_grad=grad, c_truth=M)
    # f4 FPFH + nearest features + RANSAC: a bumpy surface and its moved copy (different sampling), 3-sample RANSAC
    def bumpy(m, seed):
        r = np.random.default_rng(seed)
        u, v = r.random(m) * 0.8 - 0.4, r.random(m) * 0.8 - 0.4
        z = 0.08 * np.sin(9 * u) * np.cos(7 * v) + 0.05 * np.cos(11 * u * v) + 0.3 * u * u
        p = np.stack([u, v, z], 1)
        nrm, _ = o3r.pcd_estimate_normals(p, 20, 0.08)
        return p, o3r.pcd_orient_normals(p, nrm, (0.0, 0.0, 5.0))
    bp, bn = bumpy(700, 3)
    ap, an = bumpy(650, 4)
    G = syn.make_pose(syn.rot_z(0.5) @ syn.rot_y(0.2), (0.3, -0.1, 0.6))
    ap, an = o3r.transform_points(ap, np.linalg.inv(G)), an @ G[:3, :3]      # source = target surface seen from another frame
    fa, fb = o3r.compute_fpfh(ap, an, 60, 0.12), o3r.compute_fpfh(bp, bn, 60, 0.12)
    ij = o3r.match_features(fa, fb)
    corres = np.stack([np.arange(len(ap), dtype=np.int32), ij], 1)
    RT, rfit, rrmse, rrep = o3r.ransac_correspondence(ap, bp, corres, 0.05, ransac_n=3, edge_thr=0.9, dist_thr=0.03, max_iteration=50000, confidence=0.999, seed=5)
    out.update(g_src=ap, g_src_nrm=an, g_tgt=bp, g_tgt_nrm=bn, g_fpfh_src=fa, g_fpfh_tgt=fb, g_match=ij, g_T=RT, g_fit=np.array([rfit, rrmse]),
               g_report=np.array(rrep, np.int64), g_truth=G)
    path = os.path.join(ROOT, "tests", "golden", "hotpath_v3.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes; pose graph kept", len(oe), "of", len(edges), "edges, iterations", r1, r2,
          "| coloured ICP", it, "iterations fit", fit, "| RANSAC", rrep, "fit", rfit, "err", np.abs(RT - G).max())


if __name__ == "__main__":
    main()
