"""One GPU standing in for rank 0 of W (W = 1, 2, 4, 8): times the bench's device-resident step and the two TSDF
kernels per frame for each K2 work quantum (B2R_FUSE_XL = 4 | 2 x-layers per item).  Integration has no collective,
so a rank's time is what one GPU measures alone; this is how the quantum per world size was chosen without
spending 8 GPU-boxes per trial.  Usage: python tools/shard_probe.py [--frames 300] > profiles/rN_shard_probe.json"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import azurekinect3dreconstruction_b200 as o3d  # noqa: E402
from azurekinect3dreconstruction_b200 import _lib  # noqa: E402
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=300)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--worlds", default="1,2,4,8")
    ap.add_argument("--xls", default="4,2")
    ap.add_argument("--batches", default="16")
    ap.add_argument("--rings", default="6", help="B2R_FUSE_RING_ROUNDS values")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    depth, rgb, E, intr = bench.make_frames(args.frames, dev)
    _, H, W = depth.shape
    K = o3d.camera.PinholeCameraIntrinsic(W, H, intr["fx"], intr["fy"], intr["cx"], intr["cy"])
    TT = o3d.pipelines.integration.TSDFVolumeColorType
    lib = _lib.load()
    rows = []
    for world in [int(x) for x in args.worlds.split(",")]:
        for xl in [int(x) for x in args.xls.split(",")]:
            for batch, rr in [(int(x), int(y)) for x in args.batches.split(",") for y in args.rings.split(",")]:
                os.environ["B2R_FUSE_XL"] = str(xl)
                os.environ["B2R_FUSE_RING_ROUNDS"] = str(rr)
                vol = o3d.pipelines.integration.ScalableTSDFVolume(bench.VOXEL, bench.TRUNC, TT.RGB8, capacity_units=40000, rank=0, world=world)
                vol.set_batch(batch)

                def step():
                    vol.reset()
                    vol.integrate_raw(depth, rgb, K, E, bench.DEPTH_SCALE, bench.DEPTH_TRUNC)
                    vol.flush()
                for _ in range(3):
                    step()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.steps):
                    step()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / args.steps
                _lib.check(lib.b2r_tsdf_profile(vol._h, 2 * args.frames + 8))
                vol.reset()
                vol.integrate_raw(depth, rgb, K, E, bench.DEPTH_SCALE, bench.DEPTH_TRUNC)
                prof = (C.c_double * 5)()
                _lib.check(lib.b2r_tsdf_profile_read(vol._h, prof, _lib.stream_ptr()))
                _lib.check(lib.b2r_tsdf_profile(vol._h, 0))
                st = vol.stats()
                rows.append({"world": world, "fuse_xl": xl, "ring_rounds": rr, "batch": batch, "us_per_frame": 1e3 * ms / args.frames,
                             "rank_frames_per_s": args.frames / (ms * 1e-3), "k1_us_per_frame": 1e3 * prof[1] / max(1.0, prof[0]),
                             "k2_us_per_frame": 1e3 * prof[2] / max(1.0, prof[0]), "units": st["num_units"], "updated_voxels": st["updated_voxels"]})
                print(json.dumps(rows[-1]), file=sys.stderr, flush=True)
                del vol
    print(json.dumps({"tool": "shard_probe", "frames": args.frames, "rows": rows}))


if __name__ == "__main__":
    main()
