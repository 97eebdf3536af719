"""o3d.pipelines.integration look-alike: ScalableTSDFVolume on libb2recon.so.

Reference call sites: GoodColorReconstruction.py:111-115 (ctor), :270 (integrate), :285,305 (extract);
bundlefusion_reconstruction.py:108-112,276,309-320; capture_frames.py:18-23,157; WORKSkinda.py:533-570.
Upstream semantics: SURVEY.md Appendix A.4 / A.5.
"""
from __future__ import annotations

import ctypes as C
import enum
import os

import numpy as np
import torch

from .. import _lib
from ..camera import PinholeCameraIntrinsic
from ..geometry import PointCloud, RGBDImage, TriangleMesh
from ..utility import Vector3dVector, Vector3iVector


class TSDFVolumeColorType(enum.IntEnum):
    NoColor = 0
    RGB8 = 1
    Gray32 = 2


def _default_capacity(color_type: int) -> int:
    env = os.environ.get("B2R_TSDF_CAPACITY_UNITS")
    if env:
        return int(env)
    return 65536 if color_type else 131072   # 5.0 GiB (RGB8) / 4.0 GiB (NoColor) of zeroed unit pool


class ScalableTSDFVolume:
    """ScalableTSDFVolume(voxel_length, sdf_trunc, color_type, volume_unit_resolution=16, depth_sampling_stride=4).

    Build extensions (keyword-only, ignored by reference scripts): ``capacity_units`` (size of the unit
    pool), ``rank``/``world`` (this handle keeps only the units whose key hashes to ``rank``).
    """

    def __init__(self, voxel_length: float, sdf_trunc: float, color_type, volume_unit_resolution: int = 16,
                 depth_sampling_stride: int = 4, *, capacity_units: int = None, rank: int = 0, world: int = 1):
        _lib.require_cuda()
        self._lib = _lib.load()
        self.voxel_length = float(voxel_length)
        self.sdf_trunc = float(sdf_trunc)
        self.color_type = TSDFVolumeColorType(int(color_type))
        self.volume_unit_resolution = int(volume_unit_resolution)
        self.depth_sampling_stride = int(depth_sampling_stride)
        cap = int(capacity_units) if capacity_units else _default_capacity(int(self.color_type))
        self._h = self._lib.b2r_tsdf_create(self.voxel_length, self.sdf_trunc, int(self.color_type), self.volume_unit_resolution,
                                            self.depth_sampling_stride, cap, int(rank), int(world))
        if not self._h:
            raise RuntimeError("ScalableTSDFVolume: " + _lib.last_error())
        self.rank, self.world = int(rank), int(world)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.b2r_tsdf_destroy(h)
            except Exception:
                pass

    # ---- integrate ----------------------------------------------------------------------------------
    def integrate(self, image: RGBDImage, intrinsic: PinholeCameraIntrinsic, extrinsic) -> None:
        """ScalableTSDFVolume::Integrate. extrinsic: world -> camera, 4x4 f64."""
        d = image.depth.tensor
        c = image.color.tensor
        bad = (d.dim() != 2 or d.dtype != torch.float32 or d.shape[1] != intrinsic.width or d.shape[0] != intrinsic.height)
        color = gray = None
        cch = 0
        if self.color_type == TSDFVolumeColorType.RGB8:
            bad = bad or not (c.dtype == torch.uint8 and c.dim() == 3 and c.shape[2] == 3 and tuple(c.shape[:2]) == tuple(d.shape))
            color, cch = c, 3
        elif self.color_type == TSDFVolumeColorType.Gray32:
            bad = bad or not (c.dtype == torch.float32 and c.dim() == 2 and tuple(c.shape) == tuple(d.shape))
            gray = c
        if bad:
            raise RuntimeError("[ScalableTSDFVolume::Integrate] Unsupported image format.")
        fx, fy, cx, cy = intrinsic.fxfycxcy()
        rc = self._lib.b2r_tsdf_integrate(self._h, _lib.ptr(d), 0, 1.0, 0.0, _lib.ptr(color), cch, 0, _lib.ptr(gray),
                                          int(d.shape[1]), int(d.shape[0]), fx, fy, cx, cy, _lib.dbl16(extrinsic), _lib.stream_ptr())
        _lib.check(rc, "integrate")

    def integrate_raw(self, depth_u16: torch.Tensor, color_u8: torch.Tensor, intrinsic: PinholeCameraIntrinsic, extrinsics,
                      depth_scale: float = 1000.0, depth_trunc: float = 3.0, color_bgr: bool = False) -> None:
        """Batched fused path: raw u16 depth (B,H,W) + u8 colour (B,H,W,3|4) with the RGBD conversion fused
        (the reset + re-integrate loop of bundlefusion_reconstruction.py:309-315). Device or pinned-host tensors."""
        if depth_u16.dim() == 2:
            depth_u16, color_u8 = depth_u16[None], (color_u8[None] if color_u8 is not None else None)
        # upstream's format check, applied to the raw batch: anything else would make K1 read out of bounds
        bad = depth_u16.dim() != 3 or depth_u16.dtype not in (torch.uint16, torch.float32) or not depth_u16.is_contiguous() \
            or tuple(depth_u16.shape[1:]) != (intrinsic.height, intrinsic.width)
        if self.color_type == TSDFVolumeColorType.RGB8:
            bad = bad or color_u8 is None or color_u8.dtype != torch.uint8 or color_u8.dim() != 4 or color_u8.shape[3] not in (3, 4) \
                or tuple(color_u8.shape[:3]) != tuple(depth_u16.shape) or color_u8.device != depth_u16.device or not color_u8.is_contiguous()
        elif self.color_type == TSDFVolumeColorType.Gray32:
            bad = True                                   # Gray32 volumes take RGBDImages through integrate()
        else:
            color_u8 = None
        if bad:
            raise RuntimeError("[ScalableTSDFVolume::Integrate] Unsupported image format.")
        if not depth_u16.is_cuda and not depth_u16.is_pinned():
            # pageable host frames: stage through pinned memory so the copy ring stays asynchronous
            depth_u16 = depth_u16.view(torch.int16).pin_memory().view(torch.uint16) if depth_u16.dtype == torch.uint16 else depth_u16.pin_memory()
            color_u8 = color_u8.pin_memory() if color_u8 is not None else None
        B, H, W = depth_u16.shape
        E = np.ascontiguousarray(np.asarray(extrinsics, dtype=np.float64).reshape(-1))
        if E.size != 16 * B:
            raise RuntimeError("integrate_raw: need one 4x4 extrinsic per frame")
        fx, fy, cx, cy = intrinsic.fxfycxcy()
        cch = int(color_u8.shape[3]) if color_u8 is not None else 0
        is_u16 = 1 if depth_u16.dtype == torch.uint16 else 0
        fn = self._lib.b2r_tsdf_integrate_batch if depth_u16.is_cuda else self._lib.b2r_tsdf_integrate_host
        rc = fn(self._h, B, _lib.ptr(depth_u16), is_u16, float(depth_scale), float(depth_trunc), _lib.ptr(color_u8), cch,
                1 if color_bgr else 0, W, H, fx, fy, cx, cy, _lib.dbl16(E), _lib.stream_ptr())
        _lib.check(rc, "integrate_raw")

    def integrate_frames(self, depth_ptrs, color_ptrs, color_channels: int, intrinsic: PinholeCameraIntrinsic, extrinsics,
                         depth_scale: float = 1000.0, depth_trunc: float = 3.0, color_bgr: bool = False) -> None:
        """Frames that are not stored back to back (b2r_tsdf_integrate_frames): ``depth_ptrs`` / ``color_ptrs`` are sequences of
        DEVICE addresses of u16 depth / u8 colour images of the intrinsic's size, fused in sequence order (config 5: the slots
        of a per-tick all-gather buffer, tick-major).  The caller keeps the buffers alive until the stream has consumed them."""
        n = len(depth_ptrs)
        E = np.ascontiguousarray(np.asarray(extrinsics, dtype=np.float64).reshape(-1))
        if E.size != 16 * n or (color_ptrs is not None and len(color_ptrs) != n):
            raise RuntimeError("integrate_frames: need one 4x4 extrinsic (and one colour image) per frame")
        if self.color_type == TSDFVolumeColorType.RGB8 and (color_ptrs is None or color_channels not in (3, 4)):
            raise RuntimeError("[ScalableTSDFVolume::Integrate] Unsupported image format.")
        fx, fy, cx, cy = intrinsic.fxfycxcy()
        dp = (C.c_void_p * n)(*[int(p) for p in depth_ptrs])
        cp = (C.c_void_p * n)(*[int(p) for p in color_ptrs]) if color_ptrs is not None else None
        rc = self._lib.b2r_tsdf_integrate_frames(self._h, n, dp, 1, float(depth_scale), float(depth_trunc), cp, int(color_channels),
                                                 1 if color_bgr else 0, intrinsic.width, intrinsic.height, fx, fy, cx, cy,
                                                 _lib.dbl16(E), _lib.stream_ptr())
        _lib.check(rc, "integrate_frames")

    def reset(self) -> None:
        _lib.check(self._lib.b2r_tsdf_reset(self._h, _lib.stream_ptr()), "reset")

    def flush(self) -> None:
        """Fuse the frames that are packed but still pending (build extension; every read does this itself)."""
        _lib.check(self._lib.b2r_tsdf_flush(self._h, _lib.stream_ptr()), "flush")

    def set_batch(self, max_batch: int) -> None:
        """Frames fused per pass over the touched units (1..32; 1 = strictly per-call fusion)."""
        _lib.check(self._lib.b2r_tsdf_set_batch(self._h, int(max_batch), _lib.stream_ptr()), "set_batch")

    # ---- stats / parity hooks -----------------------------------------------------------------------
    def stats(self) -> dict:
        out = (C.c_int64 * 5)()
        _lib.check(self._lib.b2r_tsdf_stats(self._h, out, _lib.stream_ptr()), "stats")
        return dict(num_units=out[0], errors=out[1], last_touched=out[2], updated_voxels=out[3], touched_units=out[4])

    def num_units(self) -> int:
        return int(self.stats()["num_units"])

    def unit_keys(self) -> torch.Tensor:
        n = self.num_units()
        keys = torch.empty((n, 3), dtype=torch.int32, device=_lib.require_cuda())
        if n:
            _lib.check(self._lib.b2r_tsdf_export_keys(self._h, _lib.ptr(keys), n, _lib.stream_ptr()))
        return keys

    def export_units(self, first: int = 0, count: int = None):
        """(keys (n,3) i32, tsdf (n,4096), weight (n,4096), rgb (n,3,4096) or None), voxel index x*256+y*16+z."""
        n = self.num_units()
        count = n - first if count is None else count
        dev = _lib.require_cuda()
        tsdf = torch.empty((count, 4096), dtype=torch.float32, device=dev)
        wgt = torch.empty((count, 4096), dtype=torch.float32, device=dev)
        rgb = torch.empty((count, 3, 4096), dtype=torch.float32, device=dev) if self.color_type != TSDFVolumeColorType.NoColor else None
        if count:
            _lib.check(self._lib.b2r_tsdf_export_units(self._h, first, count, _lib.ptr(tsdf), _lib.ptr(wgt), _lib.ptr(rgb), _lib.stream_ptr()))
        return self.unit_keys()[first:first + count], tsdf, wgt, rgb

    def capacity(self) -> int:
        return int(self._lib.b2r_tsdf_capacity(self._h))

    def export_selected(self, ids: torch.Tensor):
        """Units by pool index (device int32 tensor) in export layout: (keys (n,3), tsdf (n,4096), weight (n,4096), rgb (n,3,4096) | None)."""
        dev = _lib.require_cuda()
        ids = ids.to(dev, dtype=torch.int32).contiguous()
        n = int(ids.shape[0])
        keys = torch.empty((n, 3), dtype=torch.int32, device=dev)
        tsdf = torch.empty((n, 4096), dtype=torch.float32, device=dev)
        wgt = torch.empty((n, 4096), dtype=torch.float32, device=dev)
        rgb = torch.empty((n, 3, 4096), dtype=torch.float32, device=dev) if self.color_type != TSDFVolumeColorType.NoColor else None
        if n:
            _lib.check(self._lib.b2r_tsdf_export_selected(self._h, _lib.ptr(ids), n, _lib.ptr(keys), _lib.ptr(tsdf), _lib.ptr(wgt), _lib.ptr(rgb),
                                                          _lib.stream_ptr()), "export_selected")
        return keys, tsdf, wgt, rgb

    def export_records(self, ids: torch.Tensor):
        """Units by pool index as pool-layout records: (keys (n,3) i32, units (n, planes * 4096) f32)."""
        dev = _lib.require_cuda()
        ids = ids.to(dev, dtype=torch.int32).contiguous()
        n = int(ids.shape[0])
        keys = torch.empty((n, 3), dtype=torch.int32, device=dev)
        units = torch.empty((n, self.unit_floats()), dtype=torch.float32, device=dev)
        if n:
            _lib.check(self._lib.b2r_tsdf_export_records(self._h, _lib.ptr(ids), n, _lib.ptr(keys), _lib.ptr(units), _lib.stream_ptr()), "export_records")
        return keys, units

    def unit_floats(self) -> int:
        return (2 if self.color_type == TSDFVolumeColorType.NoColor else 5) * 4096

    def pool_view(self) -> torch.Tensor:
        """The handle's unit pool as a (capacity, planes * 4096) f32 tensor sharing its memory (no copy): a collective can write unit
        records straight behind the units the handle holds; ``adopt_units`` then registers them."""
        fpu = C.c_int64(0)
        ptr = self._lib.b2r_tsdf_pool(self._h, C.byref(fpu))
        cap = self.capacity()

        class _Pool:
            __cuda_array_interface__ = {"shape": (cap, int(fpu.value)), "typestr": "<f4", "data": (int(ptr), False), "version": 2, "strides": None}
        t = torch.as_tensor(_Pool(), device=_lib.require_cuda())
        t._b2r_owner = self            # the view must not outlive the handle
        return t

    def adopt_units(self, keys: torch.Tensor, first: int) -> None:
        keys = keys.to(_lib.require_cuda(), dtype=torch.int32).contiguous()
        _lib.check(self._lib.b2r_tsdf_adopt_units(self._h, _lib.ptr(keys), int(first), int(keys.shape[0]), _lib.stream_ptr()), "adopt_units")

    def import_units(self, keys, tsdf, weight, rgb=None, sync: bool = True) -> None:
        dev = _lib.require_cuda()
        keys = (keys if torch.is_tensor(keys) else torch.as_tensor(np.asarray(keys))).to(dev, dtype=torch.int32).contiguous()
        tsdf = torch.as_tensor(tsdf, dtype=torch.float32).to(dev).contiguous()
        weight = torch.as_tensor(weight, dtype=torch.float32).to(dev).contiguous()
        if rgb is not None:
            rgb = torch.as_tensor(rgb, dtype=torch.float32).to(dev).contiguous()
        _lib.check(self._lib.b2r_tsdf_import_units(self._h, _lib.ptr(keys), keys.shape[0], _lib.ptr(tsdf), _lib.ptr(weight),
                                                    _lib.ptr(rgb), _lib.stream_ptr()), "import_units")
        if sync:
            torch.cuda.current_stream().synchronize()

    # ---- extraction ---------------------------------------------------------------------------------
    def extract_point_cloud(self, with_ids: bool = False) -> PointCloud:
        """ScalableTSDFVolume::ExtractPointCloud."""
        dev = _lib.require_cuda()
        cnt = C.c_int64(0)
        rc = self._lib.b2r_tsdf_extract_points(self._h, None, None, None, None, 0, C.byref(cnt), _lib.stream_ptr())
        if rc not in (_lib.OK, _lib.ERR_CAPACITY):
            _lib.check(rc, "extract_point_cloud")
        n = int(cnt.value)
        pcd = PointCloud()
        if n == 0:
            return pcd
        xyz = torch.empty((n, 3), dtype=torch.float64, device=dev)
        nrm = torch.empty((n, 3), dtype=torch.float64, device=dev)
        has_c = self.color_type != TSDFVolumeColorType.NoColor
        rgb = torch.empty((n, 3), dtype=torch.float64, device=dev) if has_c else None
        ids = torch.empty((n, 4), dtype=torch.int32, device=dev) if with_ids else None
        _lib.check(self._lib.b2r_tsdf_extract_points(self._h, _lib.ptr(xyz), _lib.ptr(rgb), _lib.ptr(nrm), _lib.ptr(ids), n,
                                                      C.byref(cnt), _lib.stream_ptr()), "extract_point_cloud")
        pcd.points = Vector3dVector(xyz)
        pcd.normals = Vector3dVector(nrm)
        if has_c:
            pcd.colors = Vector3dVector(rgb)
        pcd.edge_ids = ids
        return pcd

    def extract_triangle_mesh(self, with_ids: bool = False) -> TriangleMesh:
        """ScalableTSDFVolume::ExtractTriangleMesh.  The output arrays are sized from the previous extraction of this
        volume (+ 12 % slack), so a scan that extracts after every integrate (GoodColorReconstruction.py:270,285) makes ONE library
        call per mesh; only when the mesh outgrew the guess is the call repeated with the exact sizes it reported."""
        dev = _lib.require_cuda()
        nV, nF = C.c_int64(0), C.c_int64(0)
        has_c = self.color_type != TSDFVolumeColorType.NoColor
        capV, capF = getattr(self, "_mesh_cap", (0, 0))
        mesh = TriangleMesh()
        for attempt in range(2):
            vtx = torch.empty((capV, 3), dtype=torch.float64, device=dev) if capV else None
            vcol = torch.empty((capV, 3), dtype=torch.float64, device=dev) if (has_c and capV) else None
            tri = torch.empty((max(capF, 1), 3), dtype=torch.int32, device=dev) if capV else None
            ids = torch.empty((capV, 4), dtype=torch.int32, device=dev) if (with_ids and capV) else None
            rc = self._lib.b2r_tsdf_extract_mesh(self._h, _lib.ptr(vtx), _lib.ptr(vcol), _lib.ptr(tri), _lib.ptr(ids), capV, max(capF, 1) if capV else 0,
                                                 C.byref(nV), C.byref(nF), _lib.stream_ptr())
            v, f = int(nV.value), int(nF.value)
            if rc == _lib.ERR_CAPACITY and attempt == 0:
                capV, capF = v, f                                   # exact sizes; the count pass is cached, the second call only fills
                continue
            _lib.check(rc, "extract_triangle_mesh")
            break
        self._mesh_cap = (v + v // 8 + 1024, f + f // 8 + 2048) if v else (0, 0)
        if v == 0:
            return mesh
        mesh.vertices = Vector3dVector(vtx[:v])
        if has_c:
            mesh.vertex_colors = Vector3dVector(vcol[:v])
        mesh.triangles = Vector3iVector(tri[:f])
        mesh.edge_ids = ids[:v] if ids is not None else None
        return mesh

    def __repr__(self):
        return "ScalableTSDFVolume (B200) with %s." % ("no color" if self.color_type == 0 else "color")


def reintegrate_frames(volume: ScalableTSDFVolume, depth_u16, color_u8, pose_graph, intrinsic: PinholeCameraIntrinsic,
                       depth_scale: float = 1000.0, depth_trunc: float = 3.0, color_bgr: bool = False,
                       pose_is_extrinsic: bool = False, chunk: int = 64) -> None:
    """f3's GPU half: ``volume.reset()`` + every frame fused again under the optimised poses, through the batched path.

    capture_frames.py:92-105 passes ``np.linalg.inv(node.pose)`` as the extrinsic (``pose_is_extrinsic=False``);
    bundlefusion_reconstruction.py:306-315 passes ``node.pose`` itself (``pose_is_extrinsic=True``).
    depth_u16 (n,H,W) / color_u8 (n,H,W,3|4): device or pinned-host tensors, frame i belongs to node i."""
    n = len(pose_graph.nodes)
    if depth_u16.shape[0] < n:
        raise RuntimeError("reintegrate_frames: fewer frames than pose-graph nodes")
    volume.reset()
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        E = np.stack([nd.pose if pose_is_extrinsic else np.linalg.inv(nd.pose) for nd in pose_graph.nodes[a:b]])
        volume.integrate_raw(depth_u16[a:b], color_u8[a:b] if color_u8 is not None else None, intrinsic, E, depth_scale, depth_trunc, color_bgr)
