import cProfile, pstats, sys, os, io, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import azurekinect3dreconstruction_b200 as o3d
import bench
dev = torch.device("cuda", 0)
depth, rgb, E, intr = bench.make_frames(60, dev, 300)
K = o3d.camera.PinholeCameraIntrinsic(intr["width"], intr["height"], intr["fx"], intr["fy"], intr["cx"], intr["cy"])
dn = [np.array(depth[i].cpu().view(torch.int16).numpy().view(np.uint16)) for i in range(60)]
cn = [np.array(rgb[i].cpu().numpy()) for i in range(60)]
TT = o3d.pipelines.integration.TSDFVolumeColorType
vol = o3d.pipelines.integration.ScalableTSDFVolume(0.004, 0.02, TT.RGB8, capacity_units=16000)
def loop():
    vol.reset()
    for i in range(60):
        rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(o3d.geometry.Image(cn[i]), o3d.geometry.Image(dn[i]), depth_scale=1000.0, depth_trunc=3.0, convert_rgb_to_intensity=False)
        vol.integrate(rgbd, K, E[i])
    vol.flush(); torch.cuda.synchronize()
loop()
t0=time.perf_counter(); loop(); print("ms/frame", 1e3*(time.perf_counter()-t0)/60)
pr = cProfile.Profile(); pr.enable(); loop(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(22); print(s.getvalue()[:5000])
