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
