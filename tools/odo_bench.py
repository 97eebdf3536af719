#!/usr/bin/env python3
"""config 4 micro-benchmark: batched hybrid RGB-D odometry over consecutive NFOV frames (used for the ncu launch list)."""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import azurekinect3dreconstruction_b200 as o3d
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda", 0)
depth, rgb, E, intr = bench.make_frames(n + 1, dev, 300)
K = o3d.camera.PinholeCameraIntrinsic(intr["width"], intr["height"], intr["fx"], intr["fy"], intr["cx"], intr["cy"])
imgs = [o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(rgb[i]), o3d.geometry.Image(depth[i]), 1000.0, 3.0, True) for i in range(n + 1)]
opt = o3d.pipelines.odometry.OdometryOption(iteration_number_per_pyramid_level=o3d.utility.IntVector([20, 10, 5]), max_depth_diff=0.03, min_depth=0.0, max_depth=4.0)
J = o3d.pipelines.odometry.RGBDOdometryJacobianFromHybridTerm()
run = lambda: o3d.pipelines.odometry.compute_rgbd_odometry_batch(imgs[1:], imgs[:-1], K, None, J, opt, max_batch=n)
run(); torch.cuda.synchronize()
t0 = time.perf_counter(); r = run(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(json.dumps({"pairs": n, "ms_per_pair": 1e3 * dt / n, "pairs_per_s": n / dt, "all_ok": all(x[0] for x in r)}))
