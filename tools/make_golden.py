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


if __name__ == "__main__":
    main()
