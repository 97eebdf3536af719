"""Secondary records of bench.py's JSON line (`"secondary": {...}`): the north_star targets that are not the headline metric,
measured in the same driver-run process -- unprojection (configs[0]/[1]), extraction (configs[2] tail), RGB-D odometry
(configs[3]), the reference's own per-frame call pattern, and configs[4] (8 WFOV streams into one sharded 2 mm volume).
CUDA-event timing after warm-up; every record names its workload.  Imported by bench.py only."""
from __future__ import annotations

import time

import numpy as np
import torch

import azurekinect3dreconstruction_b200 as o3d
from azurekinect3dreconstruction_b200 import geometry, synthetic as syn

TT = o3d.pipelines.integration.TSDFVolumeColorType


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


def wall_time(fn, iters=3, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / iters


# ---- configs[0] / configs[1]: unprojection with the colour-to-depth gather -------------------------------------------------
def unproject_record(dev, peak, batch=256):
    """Kinect captures (native 640x576 u16 depth + 1280x720 BGRA) -> coloured f32 point clouds through the Python call with
    FRESH output tensors every call: B = 256 (throughput; 2.8 GB of outputs > L2), B = 2 (config 2: one tick of two
    cameras with their extrinsics folded in) and B = 1 (config 1 latency).  Algorithmic bytes: SURVEY 8(d)."""
    kin = syn.SyntheticKinect(seed=0, dropout=0.02, noise_sigma_mm=1.0, device=str(dev))
    poses = syn.orbit_trajectory(300)
    di, ci = kin.depth_intr, kin.color_intr
    Kd = o3d.camera.PinholeCameraIntrinsic(di["width"], di["height"], di["fx"], di["fy"], di["cx"], di["cy"])
    Kc = o3d.camera.PinholeCameraIntrinsic(ci["width"], ci["height"], ci["fx"], ci["fy"], ci["cx"], ci["cy"])
    H, W = di["height"], di["width"]
    nsrc = 8
    caps = [kin.render_pose(poses[i * 30], want_transformed_depth=False) for i in range(nsrc)]
    B = batch
    depth = torch.stack([torch.from_numpy(np.ascontiguousarray(caps[i % nsrc].depth).view(np.int16)).to(dev) for i in range(B)]).contiguous().view(torch.uint16)
    color = torch.stack([torch.as_tensor(caps[i % nsrc].color).to(dev) for i in range(B)]).contiguous()
    flip = np.diag([1.0, -1.0, -1.0, 1.0])
    T = np.stack([flip @ poses[(i * 30) % 300] for i in range(B)])

    def call(b):
        return geometry.unproject(depth[:b], Kd, depth_is_u16=True, depth_scale=1000.0, depth_trunc=3.0, color=color[:b], color_bgr=True,
                                  color_intrinsic=Kc, depth_to_color=kin.d2c, transforms=T[:b], dtype=torch.float32)
    r = call(B)
    n_valid = float(r["counts"].sum().item()) / B
    del r
    ms = ev_time(lambda: call(B), iters=5)
    algo = 2.0 * W * H + 4.0 * n_valid + 24.0 * n_valid
    ach = algo * B / (ms * 1e-3) / 1e9
    ms1 = ev_time(lambda: call(1), iters=30, warm=5)
    ms2 = ev_time(lambda: call(2), iters=30, warm=5)
    out = {"workload": "configs[0]/[1]: create_from_rgbd_image over Kinect captures (640x576 u16 depth + 1280x720 BGRA gathered to depth, "
                       "flip x extrinsic folded in), f32 xyz + rgb, fresh output tensors per call, through geometry.unproject",
           "batch": B, "ms_per_call": ms, "frames_per_s": B / (ms * 1e-3), "valid_points_per_frame": n_valid,
           "algorithmic_bytes_per_frame": algo, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
           "latency_ms_B1": ms1, "config2_two_camera_tick_ms_B2": ms2, "config2_ticks_per_s": 1e3 / ms2, "sensor_rate_hz": 30}
    del depth, color
    torch.cuda.empty_cache()
    return out


# ---- configs[2] tail: extraction of the bench volume ------------------------------------------------------------------------
def extract_record(vol, depth, rgb, K, E, peak):
    """extract_triangle_mesh / extract_point_cloud / compute_vertex_normals on the volume the bench step left behind (all frames
    of the trajectory), through the Open3D-shaped calls; `after_one_frame` = one more frame integrated before every extraction
    (the reference extracts right after integrating: GoodColorReconstruction.py:270,285-286), integrate time subtracted."""
    units = vol.num_units()
    mesh = vol.extract_triangle_mesh()
    V, F = len(mesh.vertices), len(mesh.triangles)
    ms_mesh = ev_time(lambda: vol.extract_triangle_mesh(), iters=5, warm=1)
    one = lambda: vol.integrate_raw(depth[:1], rgb[:1], K, E[:1], 1000.0, 3.0)
    ms_int = ev_time(lambda: (one(), vol.flush()), iters=5, warm=2)
    ms_changed = ev_time(lambda: (one(), vol.extract_triangle_mesh()), iters=5, warm=1) - ms_int
    pc = vol.extract_point_cloud()
    Kp = len(pc.points)
    ms_pts = ev_time(lambda: vol.extract_point_cloud(), iters=5, warm=1)
    ms_nrm = ev_time(lambda: mesh.compute_vertex_normals(), iters=5, warm=1)
    nvox = units * 4096.0
    algo_mesh = 8.0 * nvox + 24.0 * V + 24.0 * V + 12.0 * F
    algo_pts = 8.0 * nvox + 24.0 * Kp + 36.0 * Kp
    return {"workload": "extract_triangle_mesh / extract_point_cloud of the bench volume (%d units, 4 mm), f64 outputs, Python call" % units,
            "units": units, "vertices": V, "triangles": F, "points": Kp,
            "mesh_ms": ms_mesh, "mesh_ms_after_one_frame": ms_changed, "points_ms": ms_pts, "vertex_normals_ms": ms_nrm,
            "mesh_algorithmic_bytes": algo_mesh, "mesh_achieved": algo_mesh / (ms_mesh * 1e-3) / 1e9, "mesh_frac": algo_mesh / (ms_mesh * 1e-3) / 1e9 / peak,
            "mesh_frac_after_one_frame": algo_mesh / (max(ms_changed, 1e-6) * 1e-3) / 1e9 / peak,
            "points_algorithmic_bytes": algo_pts, "points_frac": algo_pts / (ms_pts * 1e-3) / 1e9 / peak, "peak": peak, "unit": "GB/s"}


# ---- configs[3]: frame-to-frame hybrid RGB-D odometry -----------------------------------------------------------------------
def odometry_record(depth, rgb, K, poses, pairs=32):
    """Consecutive frames of the bench trajectory as independent pairs (odo_init = I, RGBD_Odometry.py:26-35): intensity images,
    OdometryOption([20, 10, 5], 0.03, 0.0, 4.0), hybrid Jacobian.  One call per pair, and -- when the library has it -- the
    batched entry (all pairs in one launch sequence)."""
    n = min(pairs + 1, int(depth.shape[0]))
    mk = lambda i: o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(rgb[i]), o3d.geometry.Image(depth[i]), 1000.0, 3.0, True)
    imgs = [mk(i) for i in range(n)]
    opt = o3d.pipelines.odometry.OdometryOption(iteration_number_per_pyramid_level=o3d.utility.IntVector([20, 10, 5]), max_depth_diff=0.03,
                                                min_depth=0.0, max_depth=4.0)
    J = o3d.pipelines.odometry.RGBDOdometryJacobianFromHybridTerm()
    res = []

    def run():
        res.clear()
        for i in range(n - 1):
            res.append(o3d.pipelines.odometry.compute_rgbd_odometry(imgs[i + 1], imgs[i], K, np.eye(4), J, opt))
    ms = wall_time(run, iters=2, warm=1)
    errs = [float(np.linalg.norm(T_[:3, 3] - (np.linalg.inv(poses[i]) @ poses[i + 1])[:3, 3])) for i, (ok, T_, _) in enumerate(res)]
    out = {"workload": "configs[3]: compute_rgbd_odometry, hybrid Jacobian, [20,10,5] iterations, NFOV 640x576, %d independent consecutive pairs" % (n - 1),
           "pairs": n - 1, "ms_per_pair": ms / (n - 1), "pairs_per_s": (n - 1) / (ms * 1e-3), "all_ok": all(r[0] for r in res),
           "median_translation_error_m_vs_ground_truth": float(np.median(errs))}
    batched = getattr(o3d.pipelines.odometry, "compute_rgbd_odometry_batch", None)
    if batched is not None:
        src, tgt = imgs[1:], imgs[:-1]
        rb = []

        def run_b():
            rb.clear()
            rb.extend(batched(src, tgt, K, None, J, opt))
        msb = wall_time(run_b, iters=3, warm=1)
        same = all(a[0] == b[0] and np.array_equal(a[1], b[1]) for a, b in zip(res, rb))
        out.update({"batched_ms_per_pair": msb / (n - 1), "batched_pairs_per_s": (n - 1) / (msb * 1e-3), "batched_equals_per_pair_calls": bool(same)})
    return out


# ---- the reference's own call pattern, per frame --------------------------------------------------------------------------
def per_frame_api_record(depth_h, rgb_h, K, E, frames=60):
    """PAGEABLE numpy frames -> o3d.geometry.Image -> RGBDImage.create_from_color_and_depth -> volume.integrate(rgbd, K, E), one
    frame per call (1camera-detectsSceneChanges.py:270-291, GoodColorReconstruction.py:264-331), wall clock incl. every copy;
    alone, with extract_triangle_mesh() after every 10th frame, and after every frame."""
    n = min(frames, int(depth_h.shape[0]))
    dn = [np.array(depth_h[i].view(torch.int16).numpy().view(np.uint16)) for i in range(n)]      # pageable copies
    cn = [np.array(rgb_h[i].numpy()) for i in range(n)]
    vol = o3d.pipelines.integration.ScalableTSDFVolume(0.004, 0.02, TT.RGB8, capacity_units=16000)
    sizes = {}

    def loop(every):
        vol.reset()
        for i in range(n):
            rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(cn[i]), o3d.geometry.Image(dn[i]), depth_scale=1000.0,
                                                                      depth_trunc=3.0, convert_rgb_to_intensity=False)
            vol.integrate(rgbd, K, E[i])
            if every and (i + 1) % every == 0:
                m = vol.extract_triangle_mesh()
                sizes[every] = len(m.vertices)
        vol.flush()
    out = {"workload": "reference call pattern: pageable numpy frame -> Image -> RGBDImage -> integrate, one NFOV frame per call, %d frames, "
                       "4 mm / 2 cm; wall clock" % n, "frames": n}
    for every, name in ((0, "integrate_only"), (10, "mesh_every_10th_frame"), (1, "mesh_every_frame")):
        ms = wall_time(lambda: loop(every), iters=2, warm=1)
        out[name + "_frames_per_s"] = n / (ms * 1e-3)
        out[name + "_ms_per_frame"] = ms / n
    out["mesh_vertices_last"] = sizes.get(1)
    del vol
    return out


# ---- configs[4]: 8 WFOV streams, one sharded 2 mm volume ------------------------------------------------------------------
def config5_record(dev, rank, world, ticks=64, n_streams=8, ticks_per_batch=4):
    """Stream s is rendered on rank s // (8 / world) ONLY and stays in that rank's HBM as frame records.  Timed step = reset + every
    tick's frames all-gathered over NVLink (one NCCL call per batch of ticks, inside the timed region) + integrated tick-major
    into the rank's shard (unit-key hash) + flush.  Then the collective mesh extraction.  Device-timed, max over ranks."""
    import torch.distributed as dist
    from azurekinect3dreconstruction_b200.pipelines import sharded
    intr = syn.WFOV_DEPTH
    K = o3d.camera.PinholeCameraIntrinsic(intr["width"], intr["height"], intr["fx"], intr["fy"], intr["cx"], intr["cy"])
    H, W = intr["height"], intr["width"]
    ex = sharded.StreamTickExchange((H, W), 3, n_streams, ticks_per_batch=ticks_per_batch, device=dev)
    cams = syn.ring_cameras(n_streams, radius=1.5)
    mine = [s for s in range(n_streams) if s // ex.s_local == rank]
    # one source per STREAM, seeded by the stream: a camera's noise / dropout sequence is the same whichever rank renders it, so the
    # job fuses identical frames at every N (units_total / updated_voxels_total must then agree across N)
    kins = {s: syn.SyntheticKinect(depth_intrinsics=intr, seed=100 + s, dropout=0.02, noise_sigma_mm=1.0, device=str(dev), max_range_m=syn.WFOV_MAX_RANGE_M)
            for s in mine}
    packed = torch.zeros((ticks, ex.s_local, ex.rec), dtype=torch.uint8, device=dev)
    t_r0 = time.perf_counter()
    for t in range(ticks):
        drift = syn.make_pose(syn.rot_y(0.002 * t), (0.0, 0.001 * t, 0.0))      # the rig drifts: consecutive ticks are different frames
        fr = [kins[s].render_depth_aligned_rgb(drift @ cams[s]) for s in mine]
        ex.pack(torch.stack([f[0].view(torch.int16).to(dev) for f in fr]).view(torch.uint16), torch.stack([f[1].to(dev) for f in fr]),
                np.stack([np.linalg.inv(drift @ cams[s]) for s in mine]), out=packed[t])
    torch.cuda.synchronize()
    render_s = time.perf_counter() - t_r0
    cap = 120000 // world + 12000
    vol = sharded.ShardedTSDFVolume(0.002, 0.01, TT.RGB8, capacity_units=cap) if world > 1 else None
    local = vol.local if vol is not None else o3d.pipelines.integration.ScalableTSDFVolume(0.002, 0.01, TT.RGB8, capacity_units=cap)
    trunc_m = syn.WFOV_MAX_RANGE_M + 0.5

    def step():
        local.reset()
        sharded.integrate_stream_ticks(local, ex, packed, K, 1000.0, trunc_m)
        local.flush()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    step()
    e1.record()
    barrier()
    red = lambda v, op: v if world == 1 else _reduce(v, op, dev)
    ms = red(e0.elapsed_time(e1), "max")
    st = local.stats()
    units = int(red(float(st["num_units"]), "sum"))
    upd = int(red(float(st["updated_voxels"]), "sum"))
    # collective mesh extraction, timed twice (the second call is the steady state: scratch allocated)
    def extract():
        return vol.extract_triangle_mesh() if vol is not None else local.extract_triangle_mesh()
    mesh = extract()
    barrier()
    runs = []
    for _ in range(3):
        t0 = time.perf_counter()
        mesh = extract()
        barrier()
        runs.append(red(time.perf_counter() - t0, "max"))
    t_mesh = min(runs)
    phases = None
    if vol is not None:                                    # one more extraction with a device sync after every phase: where the time goes
        vol.timing = {}
        mesh = extract()
        barrier()
        phases = {k: round(1e3 * v, 3) for k, v in sorted(vol.timing.items())}
        vol.timing = None
    nfr = ticks * n_streams
    algo = 40.0 * upd + 7.0 * W * H * nfr
    out = {"workload": "configs[4]: %d WFOV 1024x1024 streams x %d ticks, voxel 2 mm, trunc 1 cm, ONE volume sharded by unit-key hash over %d GPU(s); "
                       "stream s lives on rank s // %d, per-batch NVLink all-gather of {depth, colour, pose} records inside the timed region"
                       % (n_streams, ticks, world, ex.s_local),
           "n_gpus": world, "ticks": ticks, "streams": n_streams, "ticks_per_gather": ticks_per_batch,
           "stream_frames_per_s": nfr / (ms * 1e-3), "ticks_per_s": ticks / (ms * 1e-3), "ms_per_step": ms,
           "nvlink_gathered_bytes_per_rank_per_step": ex.bytes_gathered // 2 if world > 1 else 0,
           "units_total": units, "updated_voxels_total": upd, "algorithmic_bytes_per_step": algo,
           "achieved_aggregate": algo / (ms * 1e-3) / 1e9, "unit": "GB/s",
           "mesh_gather_s": t_mesh, "mesh_gather_s_runs": runs, "mesh_phases_ms_rank0": phases, "vertices": len(mesh.vertices), "triangles": len(mesh.triangles),
           "extraction_scratch_bytes_rank0": getattr(vol, "last_extract_scratch_bytes", None) if vol is not None else None,
           "render_s": render_s}
    del packed, local, vol, ex, kins
    torch.cuda.empty_cache()
    return out


def _reduce(v, op, dev):
    import torch.distributed as dist
    t = torch.tensor([float(v)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return float(t.item())
