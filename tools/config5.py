#!/usr/bin/env python3
"""BASELINE.json configs[4]: 8 WFOV 1024x1024 streams fused at 2 mm voxels (trunc 1 cm) into ONE volume sharded by
unit-key hash over the ranks (torchrun, one process per GPU), then the collective mesh gather.

    python -m torch.distributed.run --nproc-per-node N tools/config5.py [--ticks 8] [--streams 8]

Every rank holds every stream's frames (rendered locally from the same seed: the stand-in for the per-tick frame
broadcast), integrates only the units it owns, and the ranks extract the mesh collectively.  Prints one JSON line
on rank 0: stream-frames/s for integration (device-timed, max over ranks), mesh gather time, mesh size."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import azurekinect3dreconstruction_b200 as o3d  # noqa: E402
from azurekinect3dreconstruction_b200 import synthetic as syn  # noqa: E402
from azurekinect3dreconstruction_b200.pipelines import sharded  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ticks", type=int, default=8)
    ap.add_argument("--streams", type=int, default=8)
    ap.add_argument("--capacity", type=int, default=0)
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    intr = syn.WFOV_DEPTH
    kin = syn.SyntheticKinect(depth_intrinsics=intr, seed=0, dropout=0.02, noise_sigma_mm=1.0, device=str(dev), max_range_m=syn.WFOV_MAX_RANGE_M)
    K = o3d.camera.PinholeCameraIntrinsic(intr["width"], intr["height"], intr["fx"], intr["fy"], intr["cx"], intr["cy"])
    cams = syn.ring_cameras(a.streams, radius=1.5)
    # per tick the rig drifts a little so that consecutive ticks are not identical frames
    depth, rgb, E = [], [], []
    for t in range(a.ticks):
        drift = syn.make_pose(syn.rot_y(0.002 * t), (0.0, 0.001 * t, 0.0))
        for c in cams:
            pose = drift @ c
            d, col = kin.render_depth_aligned_rgb(pose)
            depth.append(d.view(torch.int16).to(dev)); rgb.append(col.to(dev)); E.append(np.linalg.inv(pose))
    depth = torch.stack(depth).contiguous().view(torch.uint16)
    rgb = torch.stack(rgb).contiguous()
    E = np.stack(E)
    TT = o3d.pipelines.integration.TSDFVolumeColorType
    cap = a.capacity or (260000 // world + 20000)
    vol = sharded.ShardedTSDFVolume(0.002, 0.01, TT.RGB8, capacity_units=cap)

    def step():
        vol.reset()
        vol.integrate_raw(depth, rgb, K, E, 1000.0, syn.WFOV_MAX_RANGE_M + 0.5)
        vol.local.flush()
    step()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); step(); e1.record()
    dist.barrier(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    st = vol.local.stats()
    units = torch.tensor([st["num_units"]], dtype=torch.int64, device=dev)
    dist.all_reduce(units)
    t0 = time.perf_counter()
    mesh = vol.extract_triangle_mesh()
    torch.cuda.synchronize(); dist.barrier()
    t_mesh = time.perf_counter() - t0
    if rank == 0:
        nfr = a.ticks * a.streams
        print(json.dumps({"config": "configs[4]: %d WFOV 1024x1024 streams x %d ticks, voxel 2 mm, trunc 1 cm, one volume sharded over %d GPU(s)" % (a.streams, a.ticks, world),
                          "n_gpus": world, "stream_frames_per_s": nfr / (float(ms.item()) * 1e-3), "ticks_per_s": a.ticks / (float(ms.item()) * 1e-3),
                          "ms_per_step": float(ms.item()), "units_total": int(units.item()), "updated_voxels_rank0": st["updated_voxels"],
                          "mesh_gather_s": t_mesh, "vertices": len(mesh.vertices), "triangles": len(mesh.triangles)}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
