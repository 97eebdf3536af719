"""azurekinect3dreconstruction_b200 -- B200-native RGB-D reconstruction core.

A drop-in for the Open3D-shaped API the reference scripts call (``import open3d as o3d``):

    import azurekinect3dreconstruction_b200 as o3d
    rgbd = o3d.geometry.RGBDImage.create_from_color_and_depth(color, depth, convert_rgb_to_intensity=False)
    pcd  = o3d.geometry.PointCloud.create_from_rgbd_image(rgbd, intrinsic)
    vol  = o3d.pipelines.integration.ScalableTSDFVolume(0.004, 0.02, o3d.pipelines.integration.TSDFVolumeColorType.RGB8)
    vol.integrate(rgbd, intrinsic, np.linalg.inv(pose)); mesh = vol.extract_triangle_mesh()

All arithmetic runs in hand-written sm_100a CUDA behind the C-ABI of include/b2recon.h
(libb2recon.so); there is no CPU fallback.  See DESIGN.md / INTEGRATION.md.
"""
from . import camera, core, geometry, io, pipelines, synthetic, utility  # noqa: F401
from ._lib import LIB_PATH, load as load_library  # noqa: F401

__version__ = "0.1.0"
__all__ = ["camera", "core", "geometry", "io", "pipelines", "synthetic", "utility", "load_library", "LIB_PATH"]
