"""Shared test helpers."""
import math

import numpy as np


def sphere_volume(o3r, r=0.1, vl=0.01, color_type=0):
    """Analytic sphere SDF written straight into units (weights 1): |p| - r, truncated at 4 voxels."""
    vol = o3r.TSDF(vl, 4 * vl, color_type)
    ul = vl * 16
    n = int(math.ceil((r + 5 * vl) / ul))
    for kx in range(-n, n):
        for ky in range(-n, n):
            for kz in range(-n, n):
                i = np.arange(16)
                x = 0.5 * vl + vl * i + kx * ul
                y = 0.5 * vl + vl * i + ky * ul
                z = 0.5 * vl + vl * i + kz * ul
                X, Y, Z = np.meshgrid(x, y, z, indexing="ij")
                sdf = np.sqrt(X * X + Y * Y + Z * Z) - r
                t = np.clip(sdf / (4 * vl), -1, 1).astype(np.float32)
                w = np.ones_like(t)
                rgb = np.stack([np.full(t.shape, 255.0 * 0.25), 128 + 100 * X / r, np.full(t.shape, 10.0)], -1).reshape(-1, 3) if color_type else None
                vol.import_unit((kx, ky, kz), t.reshape(-1), w.reshape(-1), rgb)
    return vol


def image_normals(depth_f32, K, stride=1):
    """Camera-facing normals of the organised point grid of a depth image (cross product of the forward
    differences along u and v), in the raster order of the valid pixels -- the order unprojection emits.
    Pixels without a valid forward neighbour fall back to the viewing direction.  Test-side only: normal
    estimation is outside the hot path."""
    fx, fy, cx, cy = K
    d = np.asarray(depth_f32, np.float64)[::stride, ::stride]
    H, W = d.shape
    v, u = np.mgrid[0:H, 0:W]
    X = (u * stride - cx) * d / fx
    Y = (v * stride - cy) * d / fy
    P = np.stack([X, Y, d], -1)
    du = np.zeros_like(P)
    dv = np.zeros_like(P)
    du[:, :-1] = P[:, 1:] - P[:, :-1]
    dv[:-1, :] = P[1:, :] - P[:-1, :]
    ok = (d > 0)
    oku = np.zeros_like(ok); oku[:, :-1] = ok[:, 1:] & ok[:, :-1]
    okv = np.zeros_like(ok); okv[:-1, :] = ok[1:, :] & ok[:-1, :]
    n = np.cross(du, dv)
    ln = np.linalg.norm(n, axis=-1)
    good = oku & okv & (ln > 1e-12) & (np.abs(du[..., 2]) < 0.05) & (np.abs(dv[..., 2]) < 0.05)
    view = -P / np.maximum(np.linalg.norm(P, axis=-1, keepdims=True), 1e-12)
    n = np.where(good[..., None], n / np.maximum(ln, 1e-12)[..., None], view)
    flip = (n * view).sum(-1) < 0
    n[flip] = -n[flip]
    return np.ascontiguousarray(n[ok])
