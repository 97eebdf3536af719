#!/usr/bin/env python3
"""bench.py -- BASELINE.json metric: TSDF-integrated depth frames/sec (640x576) on 1/2/4/8 B200 + % HBM roofline.

Workload (BASELINE.json configs[2], SURVEY.md 8(d) "C3"): ScalableTSDFVolume integration of a 300-frame
synthetic orbit trajectory, NFOV 640x576 u16 depth + aligned u8 RGB, 4 mm voxels, sdf_trunc 2 cm, 16^3
units, depth_sampling_stride 4.  One STEP = volume.reset() + integrate all frames of the trajectory.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--frames F] [--impl reference]

* value   : frames/s with the frames resident in HBM (device-timed, CUDA events, max over ranks).
* e2e     : the same through the public Python API with HOST (pinned) frames: H2D copies inside the timed
            region (overlapped by the library's 3-deep ring) and a D2H read of the volume stats per step.
* roofline: tsdf_fuse_kernel, algorithmic bytes (40 B x updated voxels + 7 B x W x H per frame)
            x frames per launch over its CUDA-event duration, against MEASURED_PEAKS.json hbm_gbs; the same
            for strictly per-frame launches (batch 1) under "single_frame_pass".
* cpu_baseline: the CPU oracle (port of Open3D 0.19 legacy semantics; Open3D itself cannot be installed
            here) on the SAME 300-frame step, all host threads, one timed pass; `upstream_structure` = the same code
            parallelised as upstream is (OpenMP over x inside one unit) on the first frames of the trajectory.
* parity_checked: after that pass the GPU volume of the same frames is compared with the oracle's (unit-key set,
            TSDF + weight bit-exact, colour <= 1e-3, updated-voxel / (unit, frame) counters, mesh topology).
* --impl reference: the CPU arm on the same config: every step = reset + all 300 frames.
* N > 1   : the volume is sharded by unit-key hash; every rank sees every frame (resident in its
            HBM) and integrates only its own units; no data-path collective; strong scaling.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

VOXEL, TRUNC = 0.004, 0.02
DEPTH_SCALE, DEPTH_TRUNC = 1000.0, 3.0
RECORD_BYTES = 20          # tsdf f32 + weight f32 + rgb 3 x f32 (SURVEY 8(d) canonical record R)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--frames", type=int, default=300)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--upstream-frames", type=int, default=40, help="frames of the upstream-structure (OpenMP inside a unit) CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--c5-ticks", type=int, default=64, help="ticks of the config-5 secondary record (8 WFOV streams per tick)")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary records (configs 1, 2, 4, 5, per-frame API)")
    return ap.parse_args()


def workload_config(args):
    """Identical in both arms (b200 / reference) for the same command line: it names the workload, nothing else."""
    return {"workload": "configs[2]: ScalableTSDFVolume.integrate over a %d-frame synthetic orbit, NFOV 640x576 u16 depth + "
                        "aligned u8 RGB, voxel 4 mm, sdf_trunc 2 cm, unit 16^3, stride 4; step = reset + all frames" % args.frames,
            "frames_per_step": args.frames, "voxel_length": VOXEL, "sdf_trunc": TRUNC, "color_type": "RGB8",
            "depth_scale": DEPTH_SCALE, "depth_trunc": DEPTH_TRUNC,
            "l2_policy": "inputs (%.0f MB of frames) and the touched voxel pool are larger than the 126 MB L2" % (args.frames * 1.84)}


def make_frames(n_frames, device, trajectory_len=None):
    """Render the trajectory (torch, on `device`). Returns depth (n,H,W) uint16, rgb (n,H,W,3) uint8, extrinsics (n,4,4)."""
    import torch
    from azurekinect3dreconstruction_b200 import synthetic as syn
    kin = syn.SyntheticKinect(seed=0, dropout=0.02, noise_sigma_mm=1.0, device=str(device))
    poses = syn.orbit_trajectory(trajectory_len or n_frames)[:n_frames]
    H, W = kin.depth_intr["height"], kin.depth_intr["width"]
    depth = torch.empty((n_frames, H, W), dtype=torch.int16, device=device)
    rgb = torch.empty((n_frames, H, W, 3), dtype=torch.uint8, device=device)
    for i, p in enumerate(poses):
        d, c = kin.render_depth_aligned_rgb(p)
        depth[i] = d.view(torch.int16)
        rgb[i] = c
    E = np.stack([np.linalg.inv(p) for p in poses])
    return depth.view(torch.uint16), rgb, E, kin.depth_intr


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_reference(depth_np, rgb_np, E, intr, n_frames, steps, warmup, parallel_units=True, keep=False):
    """Times the CPU oracle (Open3D-legacy port, all host threads) on reset + n_frames frames per step.
    parallel_units=True: OpenMP over volume units; False: upstream's structure (OpenMP over x inside one unit).
    Returns (frames/s, ms/step, cores[, the volume of the last step])."""
    from oracle import o3r
    o3r.set_num_threads(os.cpu_count() or 1)       # all host cores, also under torchrun (which exports OMP_NUM_THREADS=1)
    Kt = (intr["fx"], intr["fy"], intr["cx"], intr["cy"])
    vol = o3r.TSDF(VOXEL, TRUNC, 1, 16, 4, parallel_units=parallel_units)
    times = []
    for s in range(warmup + steps):
        vol.reset()
        t0 = time.perf_counter()
        for i in range(n_frames):
            # RGBDImage conversion is part of the path the GPU arm fuses (a1), so it is timed here too
            d = o3r.depth_to_float(depth_np[i], DEPTH_SCALE, DEPTH_TRUNC)
            vol.integrate(d, Kt, E[i], rgb=rgb_np[i])
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
    total = sum(times)
    out = (n_frames * len(times) / total, 1e3 * total / len(times), o3r.num_threads())
    return out + (vol,) if keep else out


def upstream_structure_sample(depth_np, rgb_np, E, intr, nf):
    fps, ms, cores = cpu_reference(depth_np[:nf], rgb_np[:nf], E[:nf], intr, nf, 1, 0, parallel_units=False)
    return {"value": fps, "unit": "frames/s", "cores": cores,
            "sample": "reset + first %d frames of the trajectory, OpenMP over x inside one unit (upstream's loop structure), 1 pass" % nf}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  Open3D 0.19 cannot be installed or
    compiled here (no sources under /root/reference, no network), so this is the oracle port, all host threads,
    on the SAME config as the b200 arm: every step = reset + all frames of the trajectory."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    # the synthetic frames are rendered wherever it is fastest (input generation is not part of the timed path)
    rdev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))) if torch.cuda.is_available() else torch.device("cpu")
    depth, rgb, E, intr = make_frames(args.frames, rdev)
    depth_np = depth.cpu().view(torch.int16).numpy().view(np.uint16)
    rgb_np = rgb.cpu().numpy()
    fps, ms, cores = cpu_reference(depth_np, rgb_np, E, intr, args.frames, args.steps, args.warmup)
    up = upstream_structure_sample(depth_np, rgb_np, E, intr, max(1, min(args.upstream_frames, args.frames)))
    line = {"metric": "tsdf_integrated_depth_frames_per_sec_640x576", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": workload_config(args),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                             "sample": "the whole step: reset + all %d frames, OpenMP over volume units" % args.frames,
                             "upstream_structure": up},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    import azurekinect3dreconstruction_b200 as o3d
    from azurekinect3dreconstruction_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world

    depth, rgb, E, intr = make_frames(args.frames, dev)
    F, H, W = depth.shape
    K = o3d.camera.PinholeCameraIntrinsic(W, H, intr["fx"], intr["fy"], intr["cx"], intr["cy"])
    TT = o3d.pipelines.integration.TSDFVolumeColorType
    cap = int(os.environ.get("B2R_BENCH_CAPACITY_UNITS", "40000"))
    vol = o3d.pipelines.integration.ScalableTSDFVolume(VOXEL, TRUNC, TT.RGB8, capacity_units=cap, rank=rank, world=world)
    lib = _lib.load()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        vol.reset()
        vol.integrate_raw(depth, rgb, K, E, DEPTH_SCALE, DEPTH_TRUNC)
        vol.flush()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- device-resident arm ------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step_device()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms = timed(step_device, args.steps)
    clocks = sampler.stop() if sampler else None
    st = vol.stats()
    fps = args.frames * args.steps / (ms * 1e-3)

    # ---- per-kernel pass for the roofline (same workload, CUDA events around each kernel launch) --------
    peak, peak_src = hbm_peak()

    def kernel_pass(batch):
        """One step with events around every launch. Returns per-launch figures of the fusion kernel."""
        vol.set_batch(batch)
        step_device()                                  # warm (texel buffers for this batch size exist)
        _lib.check(lib.b2r_tsdf_profile(vol._h, 2 * args.frames + 8))
        vol.reset()
        vol.integrate_raw(depth, rgb, K, E, DEPTH_SCALE, DEPTH_TRUNC)
        prof = (C.c_double * 5)()
        _lib.check(lib.b2r_tsdf_profile_read(vol._h, prof, _lib.stream_ptr()))
        _lib.check(lib.b2r_tsdf_profile(vol._h, 0))
        s1 = vol.stats()
        n_upd = s1["updated_voxels"]                    # counters were zeroed by reset()
        k2_launches = max(1.0, prof[4])
        fuse_ms = prof[2] / k2_launches
        # 2 x 20 B per updated voxel + 7 B per pixel, per frame; x frames per launch
        algo = (RECORD_BYTES * 2.0 * n_upd + 7.0 * W * H * args.frames) / k2_launches
        ach = algo / (fuse_ms * 1e-3) / 1e9
        return {"frames_per_launch": args.frames / k2_launches, "algorithmic_bytes_per_launch": algo, "avg_launch_ms": fuse_ms,
                "achieved": ach, "frac": ach / peak, "fuse_kernel_launches": int(prof[4]), "pack_alloc_kernel_launches": int(prof[3]),
                "pack_alloc_kernel_ms_per_frame": prof[1] / max(1.0, prof[0]), "fuse_kernel_ms_per_frame": prof[2] / max(1.0, prof[0]),
                "updated_voxels_per_frame": n_upd / args.frames, "touched_units_per_frame": s1["touched_units"] / args.frames,
                "fuse_ms_total": prof[2], "pack_ms_total": prof[1]}

    batch = int(os.environ.get("B2R_TSDF_BATCH", "32"))
    kp = kernel_pass(batch)
    k1 = kernel_pass(1)                                 # strictly per-frame fusion: the streaming (HBM-bound) form
    vol.set_batch(batch)
    # DRAM bytes per launch come from one `ncu --set full` capture of this same command; they are quoted only when the capture
    # matches this run (GPU count, frames per launch), else null
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "fuse_kernel_traffic.json")
    if os.path.exists(tpath):
        try:
            with open(tpath) as f:
                tj = json.load(f)
            if int(tj.get("n_gpus", 1)) == world and int(tj.get("frames", 0)) == args.frames and int(tj.get("batch", 0)) == batch:
                traffic = tj.get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"kernel": "tsdf_fuse_kernel<RGB8, 2 stages, 2 x-layers per item> (batch of %d frames per pass)" % batch, "bound": "hbm", "achieved": kp["achieved"],
                "peak": peak, "unit": "GB/s", "frac": kp["frac"], "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": kp["algorithmic_bytes_per_launch"], "avg_launch_ms": kp["avg_launch_ms"],
                "frames_per_launch": kp["frames_per_launch"],
                "pack_alloc_kernel_ms_per_frame": kp["pack_alloc_kernel_ms_per_frame"], "fuse_kernel_ms_per_frame": kp["fuse_kernel_ms_per_frame"],
                "updated_voxels_per_frame": kp["updated_voxels_per_frame"], "touched_units_per_frame": kp["touched_units_per_frame"],
                "kernel_share_of_step": kp["fuse_ms_total"] / max(1e-9, ms / args.steps),
                "note": "algorithmic bytes follow SURVEY 8(d) per FRAME (every frame re-reads and re-writes its updated voxel records); "
                        "a batched pass keeps a slab on chip across its frames, so frac may exceed 1 -- DRAM traffic is in `traffic`",
                "single_frame_pass": {k: k1[k] for k in ("achieved", "frac", "avg_launch_ms", "algorithmic_bytes_per_launch",
                                                          "pack_alloc_kernel_ms_per_frame", "fuse_kernel_launches")}}
    n_launches_step = kp["fuse_kernel_launches"] + kp["pack_alloc_kernel_launches"] + 1     # + the slot-clear kernel of reset()

    # ---- end-to-end arm: host (pinned) frames through the public API -----------------------------------
    depth_h = depth.view(torch.int16).cpu().pin_memory().view(torch.uint16)
    rgb_h = rgb.cpu().pin_memory()
    d2h_bytes = 5 * 8

    # world > 1: every rank needs every frame, but PCIe carries each frame only ONCE per job -- rank r uploads its
    # 1/world share and one NCCL all-gather per super-chunk distributes it over NVLink (pipelines/sharded.py)
    exchange = None
    if world > 1:
        from azurekinect3dreconstruction_b200.pipelines import sharded
        ex_kind = os.environ.get("B2R_EXCHANGE", "peer")
        ex_args = dict(per=int(os.environ.get("B2R_EXCHANGE_PER", str(max(4, 32 // world)))), device=dev,   # one super-chunk = one 32-frame fusion pass
                       ramp=os.environ.get("B2R_EXCHANGE_RAMP", "1") != "0")
        # "peer": K1 reads the other ranks' uploaded frames over NVLink through IPC-mapped rings (no collective);
        # "nccl": one all-gather per super-chunk (the fallback, and what the CPU tests exercise over gloo)
        exchange = sharded.make_frame_exchange((H, W), 3, ex_kind, mode=os.environ.get("B2R_PEER_MODE", "pull"),
                                               slots=int(os.environ.get("B2R_EXCHANGE_SLOTS", "8")), **ex_args)

    def step_e2e():
        vol.reset()
        if exchange is None:
            vol.integrate_raw(depth_h, rgb_h, K, E, DEPTH_SCALE, DEPTH_TRUNC)
        else:
            sharded.integrate_host_frames(vol, depth_h, rgb_h, K, E, DEPTH_SCALE, DEPTH_TRUNC, exchange=exchange)
        vol.stats()            # fuses what is pending + D2H read of the step's result (units allocated / voxels updated)

    for _ in range(3):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    fps_e2e = args.frames * args.steps / (ms_e2e * 1e-3)
    st_e2e = vol.stats()
    if exchange is not None and os.environ.get("B2R_E2E_TIMELINE"):
        # one more step with an event at the start and end of every upload / gather / integrate stage (rank 0's view): what
        # bounds the end-to-end arm at N > 1 is read off this timeline (profiles/r2_e2e_timeline_nN.json)
        exchange.timeline = []
        barrier()
        t_origin = torch.cuda.Event(enable_timing=True)
        t_origin.record()
        step_e2e()
        barrier()
        if rank == 0:
            rows = [{"stage": st_, "superchunk": i, "t_ms": t_origin.elapsed_time(e)} for st_, i, e in exchange.timeline]
            with open(os.environ["B2R_E2E_TIMELINE"], "w") as f:
                json.dump({"n_gpus": world, "frames": args.frames, "per": exchange.per, "ramp": exchange.ramp, "exchange": type(exchange).__name__, "events": rows}, f)
        exchange.timeline = None
    # the end-to-end arm must leave this rank's shard exactly as the device-resident arm did (same units, same voxel updates)
    if (st_e2e["num_units"], st_e2e["updated_voxels"], st_e2e["errors"]) != (st["num_units"], st["updated_voxels"], st["errors"]):
        raise RuntimeError("e2e arm diverged from the device arm: %r vs %r" % (st_e2e, st))
    h2d_bytes = int(depth_h.numel() * 2 + rgb_h.numel())

    # ---- secondary records: the other north_star targets, in the same driver-run line -------------------------------------
    secondary = None
    if not args.no_secondary:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_secondary as bs
        secondary = {}

        def record(name, fn, *a, **k):
            try:
                secondary[name] = fn(*a, **k)
            except Exception as e:                     # a secondary record never takes the headline line down with it
                secondary[name] = {"error": repr(e)[:300]}
        if rank == 0:
            poses = None
            from azurekinect3dreconstruction_b200 import synthetic as syn
            poses = syn.orbit_trajectory(args.frames)
            record("unproject", bs.unproject_record, dev, peak)
            if world == 1:
                record("extract", bs.extract_record, vol, depth, rgb, K, E, peak)
                step_device()                          # extract_record integrated extra frames: restore the step's volume
            record("odometry", bs.odometry_record, depth, rgb, K, poses)
            record("per_frame_api", bs.per_frame_api_record, depth_h, rgb_h, K, E)
        # config 5 is collective: every rank takes part
        try:
            c5 = bs.config5_record(dev, rank, world, ticks=args.c5_ticks)
        except Exception as e:
            c5 = {"error": repr(e)[:300]}
        secondary["config5"] = c5

    # ---- CPU baseline (rank 0, N = 1): the same 300-frame step on the host cores, then the parity check -----------------
    cpu = parity = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import compare
        dn = depth_h.view(torch.int16).numpy().view(np.uint16)
        cn = rgb_h.numpy()
        cfps, cms, cores, ref_vol = cpu_reference(dn, cn, E, intr, args.frames, 1, 0, keep=True)
        cpu = {"value": cfps, "unit": "frames/s", "cores": cores, "kind": "port",
               "sample": "the whole step: reset + all %d frames, 1 timed pass, OpenMP over volume units" % args.frames,
               "upstream_structure": upstream_structure_sample(dn, cn, E, intr, max(1, min(args.upstream_frames, args.frames)))}
        # the volume of the bench step (same 300 frames, batch 32) against the oracle's
        parity = {"frames": args.frames}
        parity.update(compare.volume_vs_oracle(vol, ref_vol))
        parity["bit_exact"] = bool(parity["tsdf_weight_bit_exact"] and parity["counters_equal"])
        parity["mesh"] = compare.mesh_vs_oracle(vol, ref_vol, VOXEL)
        del ref_vol

    if rank == 0:
        line = {"metric": "tsdf_integrated_depth_frames_per_sec_640x576", "value": fps, "unit": "frames/s", "n_gpus": n_gpus,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": workload_config(args), "impl": "b200",
                "sharding": {"rule": "unit-key hash over %d rank(s)" % world, "units_allocated_rank0": st["num_units"]},
                "roofline": roofline, "cpu_baseline": cpu, "parity_checked": parity, "secondary": secondary,
                "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes * world,
                        "ms_per_step": ms_e2e / args.steps,
                        "ingest": ("host frames -> one GPU over PCIe (6-deep ring of 8-frame chunks)" if world == 1 else
                                   "each rank uploads 1/%d of the frames over its PCIe link; %s (h2d_bytes_per_step is the whole job's)" % (
                                       world, "the other ranks' K1 reads them over NVLink from IPC-mapped rings, ordered by stream memory operations"
                                       if isinstance(exchange, sharded.PeerFrameExchange) else "one NCCL all-gather per super-chunk hands every rank all frames"))},
                "gpu_launches": args.steps * n_launches_step, "clocks": clocks}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def _main_with_clean_stdout():
    """The contract is ONE JSON line on stdout: libraries (NCCL prints its version banner to stdout) are kept off it by
    pointing fd 1 at stderr for the duration of the run and printing the line to the saved descriptor."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    real = os.fdopen(saved, "w")
    import builtins
    _print = builtins.print

    def print_json(*a, **k):
        k.setdefault("file", real)
        _print(*a, **k)
        real.flush()
    builtins.print = print_json
    try:
        return main()
    finally:
        builtins.print = _print
        sys.stdout.flush()


if __name__ == "__main__":
    sys.exit(_main_with_clean_stdout())
