"""Debug: per-CTA timeline of one tsdf_pack_alloc_kernel launch (needs the -DB2R_K1_TRACE variant, B2R_LIB=...)."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import azurekinect3dreconstruction_b200 as o3d
from azurekinect3dreconstruction_b200 import _lib
TT = o3d.pipelines.integration.TSDFVolumeColorType
dev = torch.device("cuda:0")
depth, rgb, E, di = bench.make_frames(64, dev, 300)
K = o3d.camera.PinholeCameraIntrinsic(di["width"], di["height"], di["fx"], di["fy"], di["cx"], di["cy"])
vol = o3d.pipelines.integration.ScalableTSDFVolume(0.004, 0.02, TT.RGB8)
for rep in range(3):
    vol.integrate_raw(depth, rgb, K, E, 1000.0, 3.0)
torch.cuda.synchronize()
lib = C.CDLL(os.environ["B2R_LIB"])
n = 8192
out = np.zeros((n, 4), dtype=np.uint64)
rc = lib.b2r_debug_k1_trace(out.ctypes.data_as(C.c_void_p), n)
print("rc", rc)
marks = np.zeros((8192, 8), dtype=np.uint64)
lib.b2r_debug_k1_marks(marks.ctypes.data_as(C.c_void_p), 8192)
al = out[:, 3] == 0
mk = marks[(out[:, 1] > 0) & al].astype(np.float64)
start = out[(out[:, 1] > 0) & al][:, 0].astype(np.float64)
end = out[(out[:, 1] > 0) & al][:, 1].astype(np.float64)
stages = {"launch_to_mark0": (mk[:, 0] - start).mean() / 1e3, "depth_load_f64_chain": (mk[:, 1] - mk[:, 0]).mean() / 1e3,
          "key_set_build": (mk[:, 2] - mk[:, 1]).mean() / 1e3, "probe_atomics_append": (mk[:, 4] - mk[:, 2]).mean() / 1e3, "exit": (end - mk[:, 4]).mean() / 1e3}
print(stages, file=sys.stderr)
out = out[out[:, 1] > 0]
t0 = out[:, 0].min()
import json
rows = []
for role in (0, 1):
    m = out[out[:, 3] == role]
    if len(m) == 0: continue
    life = (m[:, 1] - m[:, 0]).astype(np.float64) / 1e3
    rows.append({"role": "pack" if role else "alloc", "ctas": int(len(m)), "life_us_mean": float(life.mean()), "life_us_p10": float(np.percentile(life, 10)),
                 "life_us_p90": float(np.percentile(life, 90)), "first_start_us": float((m[:, 0].min() - t0) / 1e3), "last_end_us": float((m[:, 1].max() - t0) / 1e3),
                 "sum_life_us": float(life.sum())})
print(json.dumps({"kernel_span_us": float((out[:, 1].max() - t0) / 1e3), "roles": rows}, indent=1))
