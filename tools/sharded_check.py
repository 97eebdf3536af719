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
    mesh = vol.extract_triangle_mesh()                      # gather to rank 0
    held = vol.last_extract_units_held
    pcd = vol.extract_point_cloud()
    mesh_all = vol.extract_triangle_mesh(gather="all")      # every rank gets the whole mesh
    n_local = vol.local.num_units()
    ok = True
    if rank == 0:
        ref = single.extract_triangle_mesh()
        # the slab-order concatenation IS the single-GPU mesh: same vertex order, same triangle order, same colours
        ok = np.array_equal(np.asarray(mesh.vertices), np.asarray(ref.vertices)) and np.array_equal(np.asarray(mesh.triangles), np.asarray(ref.triangles)) \
            and np.array_equal(np.asarray(mesh.vertex_colors), np.asarray(ref.vertex_colors))
        rp = single.extract_point_cloud()
        ok = ok and np.array_equal(np.asarray(pcd.points), np.asarray(rp.points)) and np.array_equal(np.asarray(pcd.normals), np.asarray(rp.normals))
        ok = ok and np.array_equal(np.asarray(mesh_all.triangles), np.asarray(ref.triangles))
        print("sharded_check world=%d: units(rank0 shard)=%d of %d, held by rank 0 during extraction=%d, V=%d F=%d points=%d -> %s" % (
            world, n_local, single.num_units(), held, len(mesh.vertices), len(mesh.triangles), len(pcd.points), "OK" if ok else "MISMATCH"))
    else:
        ok = len(mesh.vertices) == 0 and len(mesh_all.vertices) > 0
    # host-frame ingest: both exchanges must leave this rank's shard exactly as direct integration of device frames does
    fr = [kin.render_depth_aligned_rgb(p) for p in syn.orbit_trajectory(21, yaw_sweep_deg=20.0)]
    dh = torch.stack([f[0].view(torch.int16) for f in fr]).contiguous().pin_memory().view(torch.uint16)
    ch = torch.stack([f[1] for f in fr]).contiguous().pin_memory()
    Es = np.stack([np.linalg.inv(p) for p in syn.orbit_trajectory(21, yaw_sweep_deg=20.0)])
    mk = lambda: o3d.pipelines.integration.ScalableTSDFVolume(0.01, 0.04, TT.RGB8, capacity_units=4096, rank=rank, world=world)
    direct = mk()
    direct.integrate_raw(dh.cuda(), ch.cuda(), K, Es, 1000.0, 3.0)
    def state(v):                                            # pool order depends on allocation races: compare in key order
        keys, t, w, c = v.export_units()
        o = torch.as_tensor(np.lexsort(keys.cpu().numpy().T[::-1]), device=t.device)
        return keys[o], t[o], w[o], c[o]
    want = state(direct)
    for kind in ("peer:pull", "peer:direct", "nccl"):
        kw = dict(per=2, device=torch.device("cuda", local), slots=3)
        ex = sharded.PeerFrameExchange((144, 160), 3, mode=kind.split(":")[1], **kw) if kind.startswith("peer") else sharded.HostFrameExchange((144, 160), 3, **kw)
        for rep in range(2):                                  # twice: ring slots and counters are reused across runs
            v = mk()
            sharded.integrate_host_frames(v, dh, ch, K, Es, 1000.0, 3.0, exchange=ex)
            got = state(v)
            same = all(torch.equal(a, b) for a, b in zip(got, want))
            ok = ok and same
            if not same:
                print("rank %d: %s exchange run %d differs from direct integration" % (rank, kind, rep))
        if kind.startswith("peer"):
            ex.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = bool(flag.item())
    if rank == 0:
        print("sharded_check world=%d: host-frame exchanges (peer pull, peer direct, nccl) vs direct integration -> %s" % (world, "OK" if ok else "MISMATCH"))
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
