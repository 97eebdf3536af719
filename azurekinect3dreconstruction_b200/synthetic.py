"""Synthetic Azure-Kinect-shaped frame source (replaces PyK4A camera I/O).

The reference pulls frames with ``k4a.get_capture()`` and reads ``capture.color`` (720x1280x4 BGRA u8),
``capture.depth`` (576x640 u16 mm) and ``capture.transformed_depth`` (720x1280 u16 mm)
(/root/reference/CodeThatKindaWorks/test_device.py:23-32, Depth2PC.py:183-197).  There is no camera
here, so an analytic scene is ray-cast with the same shapes/dtypes (SURVEY.md 8(d) "Synthetic inputs").

All math is torch (float64) so the same code renders on the CPU for tests and on cuda:0 for the bench.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np
import torch

# Intrinsics the reference declares (CodeThatWorks/BetterVisualizer2CamWorksCOLOR.py:110-127)
NFOV_DEPTH = dict(width=640, height=576, fx=504.2, fy=504.1, cx=319.8, cy=287.9)
COLOR_720P = dict(width=1280, height=720, fx=605.6, fy=605.9, cx=637.7, cy=364.3)
WFOV_DEPTH = dict(width=1024, height=1024, fx=504.2, fy=504.2, cx=511.5, cy=511.5)
NFOV_MAX_RANGE_M = 3.86
WFOV_MAX_RANGE_M = 2.21


def rot_x(a: float) -> np.ndarray:
    c, s = math.cos(a), math.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)


def rot_y(a: float) -> np.ndarray:
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)


def rot_z(a: float) -> np.ndarray:
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)


def make_pose(R: np.ndarray, t) -> np.ndarray:
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = np.asarray(t, dtype=np.float64)
    return T


def default_depth_to_color() -> np.ndarray:
    """Depth-camera -> colour-camera rigid transform (about -32 mm baseline, 6 degree tilt)."""
    return make_pose(rot_x(math.radians(6.0)), (-0.032, -0.002, 0.004))


@dataclass
class Scene:
    """Axis-aligned room (camera inside) + spheres + one tilted plane patch. y points down."""
    room_min: Tuple[float, float, float] = (-2.0, -1.3, -2.0)
    room_max: Tuple[float, float, float] = (2.0, 1.3, 2.0)
    spheres: List[Tuple[float, float, float, float]] = field(default_factory=lambda: [
        (0.9, 0.7, 1.5, 0.40), (-0.8, 0.9, 1.6, 0.30), (0.1, -0.2, 1.75, 0.25),
        (1.6, 0.6, -0.2, 0.35), (-1.6, 0.5, 0.3, 0.30)])
    # plane: point, normal, half extent (disc radius)
    plane_point: Tuple[float, float, float] = (-0.3, 0.4, 1.2)
    plane_normal: Tuple[float, float, float] = (0.35, -0.25, -0.9)
    plane_radius: float = 0.45

    def raycast(self, origin: torch.Tensor, dirs: torch.Tensor):
        """origin (3,), dirs (N,3) (not normalised). Returns t (N,) with +inf for a miss."""
        dev, dt = dirs.device, dirs.dtype
        inf = torch.full((dirs.shape[0],), float("inf"), dtype=dt, device=dev)
        t_best = inf.clone()
        # room: leaving the box through the far slab of each axis
        lo = torch.tensor(self.room_min, dtype=dt, device=dev)
        hi = torch.tensor(self.room_max, dtype=dt, device=dev)
        with torch.no_grad():
            inv = 1.0 / torch.where(dirs == 0, torch.full_like(dirs, 1e-300), dirs)
            t1 = (lo - origin) * inv
            t2 = (hi - origin) * inv
            t_exit = torch.maximum(t1, t2).min(dim=1).values
            t_best = torch.minimum(t_best, torch.where(t_exit > 0, t_exit, inf))
            dd = (dirs * dirs).sum(1)
            for (sx, sy, sz, r) in self.spheres:
                c = torch.tensor((sx, sy, sz), dtype=dt, device=dev)
                oc = origin - c
                b = (dirs * oc).sum(1)
                cc = (oc * oc).sum() - r * r
                disc = b * b - dd * cc
                ok = disc > 0
                sq = torch.sqrt(torch.clamp(disc, min=0))
                t0 = (-b - sq) / dd
                t_s = torch.where(ok & (t0 > 1e-6), t0, inf)
                t_best = torch.minimum(t_best, t_s)
            n = torch.tensor(self.plane_normal, dtype=dt, device=dev)
            n = n / n.norm()
            p0 = torch.tensor(self.plane_point, dtype=dt, device=dev)
            denom = (dirs * n).sum(1)
            tp = ((p0 - origin) * n).sum() / torch.where(denom == 0, torch.full_like(denom, 1e-300), denom)
            hit = origin + dirs * tp[:, None]
            inside = ((hit - p0) ** 2).sum(1) <= self.plane_radius ** 2
            t_p = torch.where((tp > 1e-6) & inside, tp, inf)
            t_best = torch.minimum(t_best, t_p)
        return t_best

    @staticmethod
    def color_at(p: torch.Tensor) -> torch.Tensor:
        """Procedural RGB in [0,255] (float) from world xyz: smooth field + 5 cm checker."""
        x, y, z = p[:, 0], p[:, 1], p[:, 2]
        smooth = torch.stack([0.5 + 0.5 * torch.sin(1.7 * x + 0.6 * z), 0.5 + 0.5 * torch.sin(2.3 * y + 0.4 * x + 1.0),
                              0.5 + 0.5 * torch.sin(1.1 * z - 0.9 * y + 2.0)], dim=1)
        chk = (torch.floor(x / 0.05) + torch.floor(y / 0.05) + torch.floor(z / 0.05)) % 2
        val = 0.25 + 0.55 * smooth + 0.2 * chk[:, None]
        return torch.clamp(val, 0.0, 1.0) * 255.0


def pixel_dirs(intr: dict, device, dtype=torch.float64) -> torch.Tensor:
    W, H = intr["width"], intr["height"]
    u = torch.arange(W, dtype=dtype, device=device)
    v = torch.arange(H, dtype=dtype, device=device)
    x = (u - intr["cx"]) / intr["fx"]
    y = (v - intr["cy"]) / intr["fy"]
    d = torch.stack([x[None, :].expand(H, W), y[:, None].expand(H, W), torch.ones(H, W, dtype=dtype, device=device)], dim=2)
    return d.reshape(-1, 3)


@dataclass
class Capture:
    """Mirrors the fields the reference reads off a pyk4a capture."""
    color: np.ndarray              # (Hc, Wc, 4) u8 BGRA
    depth: np.ndarray              # (Hd, Wd) u16 mm
    transformed_depth: np.ndarray  # (Hc, Wc) u16 mm, depth rendered in the colour camera
    pose: np.ndarray               # world <- depth camera, 4x4 f64 (ground truth)


class SyntheticKinect:
    """Azure-Kinect-shaped source: start()/stop()/get_capture() like pyk4a.PyK4A."""

    def __init__(self, scene: Optional[Scene] = None, depth_intrinsics: dict = NFOV_DEPTH,
                 color_intrinsics: dict = COLOR_720P, depth_to_color: Optional[np.ndarray] = None,
                 max_range_m: float = NFOV_MAX_RANGE_M, seed: int = 0, dropout: float = 0.0,
                 noise_sigma_mm: float = 0.0, device: str = "cpu", poses: Optional[List[np.ndarray]] = None):
        self.scene = scene or Scene()
        self.depth_intr = dict(depth_intrinsics)
        self.color_intr = dict(color_intrinsics)
        self.d2c = default_depth_to_color() if depth_to_color is None else np.asarray(depth_to_color, dtype=np.float64)
        self.max_range = max_range_m
        self.gen = torch.Generator(device="cpu")
        self.gen.manual_seed(seed)
        self.dropout = dropout
        self.noise = noise_sigma_mm
        self.device = torch.device(device)
        self.poses = poses if poses is not None else [np.eye(4)]
        self._i = 0
        self._dirs_d = pixel_dirs(self.depth_intr, self.device)
        self._dirs_c = pixel_dirs(self.color_intr, self.device)
        self.serial = "SYNTH%06d" % seed
        self.opened = False

    # pyk4a-like lifecycle
    def start(self):
        self.opened = True

    def stop(self):
        self.opened = False

    @property
    def calibration(self):
        return dict(depth=self.depth_intr, color=self.color_intr, depth_to_color=self.d2c)

    def _render(self, pose_wc: np.ndarray, dirs_cam: torch.Tensor, want_color: bool):
        T = torch.tensor(pose_wc, dtype=torch.float64, device=self.device)
        R, o = T[:3, :3], T[:3, 3]
        dirs_w = dirs_cam @ R.T
        t = self.scene.raycast(o, dirs_w)
        col = None
        if want_color:
            tt = torch.where(torch.isfinite(t), t, torch.zeros_like(t))
            p = o + dirs_w * tt[:, None]
            col = self.scene.color_at(p)
        return t, col

    def _depth_u16(self, t: torch.Tensor, shape) -> torch.Tensor:
        z = torch.where(torch.isfinite(t) & (t <= self.max_range), t, torch.zeros_like(t))
        mm = z * 1000.0
        if self.noise > 0:
            n = torch.randn(mm.shape, generator=self.gen, dtype=torch.float64).to(mm.device)
            mm = torch.where(mm > 0, mm + self.noise * n, mm)
        mm = torch.clamp(torch.floor(mm + 0.5), 0, 65535)
        if self.dropout > 0:
            keep = torch.rand(mm.shape, generator=self.gen, dtype=torch.float64).to(mm.device) >= self.dropout
            mm = torch.where(keep, mm, torch.zeros_like(mm))
        return mm.to(torch.int32).reshape(shape)

    def render_pose(self, pose_wd: np.ndarray, want_transformed_depth: bool = True) -> Capture:
        Hd, Wd = self.depth_intr["height"], self.depth_intr["width"]
        Hc, Wc = self.color_intr["height"], self.color_intr["width"]
        t_d, _ = self._render(pose_wd, self._dirs_d, False)
        depth = self._depth_u16(t_d, (Hd, Wd)).cpu().numpy().astype(np.uint16)
        pose_wc = pose_wd @ np.linalg.inv(self.d2c)
        t_c, col = self._render(pose_wc, self._dirs_c, True)
        rgb = torch.clamp(torch.floor(col + 0.5), 0, 255).to(torch.uint8).reshape(Hc, Wc, 3)
        bgra = torch.cat([rgb.flip(2), torch.full((Hc, Wc, 1), 255, dtype=torch.uint8, device=rgb.device)], dim=2)
        tdepth = self._depth_u16(t_c, (Hc, Wc)).cpu().numpy().astype(np.uint16) if want_transformed_depth else np.zeros((Hc, Wc), np.uint16)
        return Capture(color=bgra.cpu().numpy(), depth=depth, transformed_depth=tdepth, pose=np.array(pose_wd))

    def get_capture(self) -> Capture:
        pose = self.poses[self._i % len(self.poses)]
        self._i += 1
        return self.render_pose(pose)

    def render_depth_aligned_rgb(self, pose_wd: np.ndarray):
        """Depth u16 + RGB u8 rendered *from the depth camera* (a perfectly aligned pair on the native
        depth grid; this is what BASELINE's 640x576 metric integrates). Returns torch tensors on self.device."""
        Hd, Wd = self.depth_intr["height"], self.depth_intr["width"]
        t_d, col = self._render(pose_wd, self._dirs_d, True)
        depth = self._depth_u16(t_d, (Hd, Wd)).to(torch.uint16) if hasattr(torch, "uint16") else None
        rgb = torch.clamp(torch.floor(col + 0.5), 0, 255).to(torch.uint8).reshape(Hd, Wd, 3)
        return depth, rgb


def orbit_trajectory(n: int = 300, radius: float = 0.6, yaw_sweep_deg: float = 120.0, bob: float = 0.05) -> List[np.ndarray]:
    """Config-3 trajectory: smooth orbit, camera looking outward. world <- camera."""
    poses = []
    for i in range(n):
        s = i / max(n - 1, 1)
        th = math.radians(-0.5 * yaw_sweep_deg + yaw_sweep_deg * s)
        R = rot_y(th)
        fwd = R[:, 2]
        pos = radius * fwd + np.array([0.0, bob * math.sin(2 * math.pi * 2 * s), 0.0])
        poses.append(make_pose(R, pos))
    return poses


def ring_cameras(n: int = 8, radius: float = 1.5) -> List[np.ndarray]:
    """Config-5: n cameras on a ring looking inward at the origin."""
    poses = []
    for i in range(n):
        th = 2 * math.pi * i / n
        R = rot_y(th + math.pi)          # forward = -radial
        pos = np.array([radius * math.sin(th), 0.0, radius * math.cos(th)])
        poses.append(make_pose(R, pos))
    return poses


# ---- the reference's raw-frame layout: frames/frame_%06d.npz with keys color, depth (Depth2PC.py:190-192)
def save_frame_npz(directory: str, index: int, color: np.ndarray, depth: np.ndarray) -> str:
    os.makedirs(directory, exist_ok=True)
    path = os.path.join(directory, "frame_%06d.npz" % index)
    np.savez(path, color=color, depth=depth)
    return path


def load_frame_npz(path: str):
    with np.load(path) as z:
        return z["color"], z["depth"]
