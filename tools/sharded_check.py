#!/usr/bin/env python3
"""torchrun --nproc-per-node N tools/sharded_check.py : the real NCCL path of pipelines/sharded.py.
Every rank renders the same small trajectory, integrates its shard, the ranks extract collectively, and rank 0
compares the gathered mesh / cloud with a single-GPU volume built from the same frames (must be identical)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import azurekinect3dreconstruction_b200 as o3d  # noqa: E402
from azurekinect3dreconstruction_b200 import synthetic as syn  # noqa: E402
from azurekinect3dreconstruction_b200.pipelines import sharded  # noqa: E402


def canon(m):
    v, t = np.asarray(m.vertices), np.asarray(m.triangles)
    o = np.lexsort(v.T[::-1]); inv = np.empty_like(o); inv[o] = np.arange(len(o))
    t = inv[t]
    return v[o], t[np.lexsort(t.T[::-1])]


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    intr = dict(width=160, height=144, fx=126.05, fy=126.025, cx=79.95, cy=71.975)
    kin = syn.SyntheticKinect(depth_intrinsics=intr, seed=0, dropout=0.02)
    K = o3d.camera.PinholeCameraIntrinsic(160, 144, intr["fx"], intr["fy"], intr["cx"], intr["cy"])
    TT = o3d.pipelines.integration.TSDFVolumeColorType
    vol = sharded.ShardedTSDFVolume(0.01, 0.04, TT.RGB8, capacity_units=4096)
    single = o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, TT.RGB8, capacity_units=4096) if rank == 0 else None
    for pose in syn.orbit_trajectory(4, yaw_sweep_deg=8.0):
        d, c = kin.render_depth_aligned_rgb(pose)
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(c.numpy()), o3d.geometry.Image(d.view(torch.int16).numpy().view(np.uint16)), 1000.0, 3.0, False)
        E = np.linalg.inv(pose)
        vol.integrate(rgbd, K, E)
        if single is not None:
            single.integrate(rgbd, K, E)
    mesh = vol.extract_triangle_mesh()
    pcd = vol.extract_point_cloud()
    n_local = vol.local.num_units()
    ok = True
    if rank == 0:
        ref = single.extract_triangle_mesh()
        a, b = canon(mesh), canon(ref)
        ok = np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        rp = single.extract_point_cloud()
        ok = ok and len(pcd.points) == len(rp.points) and np.array_equal(np.sort(np.asarray(pcd.points), 0), np.sort(np.asarray(rp.points), 0))
        print("sharded_check world=%d: units(rank0)=%d of %d, V=%d F=%d points=%d -> %s" % (world, n_local, single.num_units(), len(mesh.vertices),
                                                                                        len(mesh.triangles), len(pcd.points), "OK" if ok else "MISMATCH"))
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
