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
n = 270 * 16
out = np.zeros((n, 4), dtype=np.uint64)
rc = lib.b2r_debug_k1_trace(out.ctypes.data_as(C.c_void_p), n)
print("rc", rc)
np.save("gpurun_out/k1_trace.npy", out)
