"""Checker helpers shared by tests/ and bench.py's cpu_baseline leg: a GPU volume / mesh against the CPU oracle.

TEST INFRASTRUCTURE ONLY (like everything under oracle/): never imported by the product package.  The comparisons run
on the GPU (the oracle's planes are uploaded chunk by chunk), so they finish in seconds at the BASELINE sizes
(640x576 / 4 mm: ~8 k units; 1024x1024 / 2 mm: ~30 k units = 2.4 GB of voxel records)."""
from __future__ import annotations

import numpy as np


def _lex_order(keys: np.ndarray) -> np.ndarray:
    return np.lexsort((keys[:, 2], keys[:, 1], keys[:, 0]))


def volume_vs_oracle(vol, ref, color_tol: float = 1e-3, chunk: int = 2048) -> dict:
    """vol: azurekinect3dreconstruction_b200 ScalableTSDFVolume; ref: oracle.o3r.TSDF fed the same frames.

    Bars (north_star / DESIGN.md section 1): unit-key SET equal; TSDF and weight planes BIT-exact; colour within
    `color_tol` on the 0..255 scale (named deviation: f32 running mean here, f64 upstream); the kernel's roofline
    counters equal the oracle's (updated voxel records, (unit, frame) pairs).  Raises AssertionError on any miss."""
    import torch
    keys, tsdf, wgt, rgb = vol.export_units()
    gk = keys.cpu().numpy()
    rk = ref.keys()
    assert len(gk) == len(rk), "unit count differs: %d vs oracle %d" % (len(gk), len(rk))
    go, ro = _lex_order(gk), _lex_order(rk)
    assert np.array_equal(gk[go], rk[ro]), "allocated unit-key sets differ"
    # oracle unit j  <->  GPU unit g_of_r[j]
    g_of_r = np.empty(len(rk), np.int64)
    g_of_r[ro] = go
    dev = tsdf.device
    max_dc, exact_t, exact_w, nz = 0.0, True, True, 0
    for first in range(0, len(rk), chunk):
        _, t, w, c = ref.export_range(first, chunk)
        idx = torch.as_tensor(g_of_r[first:first + len(t)], device=dev)
        tt, ww = torch.from_numpy(t).to(dev), torch.from_numpy(w).to(dev)
        exact_w = exact_w and torch.equal(wgt[idx], ww)
        # bit-exact, not just numerically equal (a -0.0 / +0.0 mix would pass ==)
        exact_t = exact_t and torch.equal(tsdf[idx].view(torch.int32), tt.view(torch.int32))
        nz += int((ww > 0).sum())
        if rgb is not None and c is not None:
            max_dc = max(max_dc, float((rgb[idx].double() - torch.from_numpy(c).to(dev)).abs().max()))
    assert exact_w, "weights differ from the oracle"
    assert exact_t, "TSDF differs from the oracle (expected bit-exact)"
    assert max_dc < color_tol, "colour differs from the oracle by %g (0..255 scale)" % max_dc
    st = vol.stats()
    assert int(st["errors"]) == 0
    assert int(st["updated_voxels"]) == ref.total_updated(), "updated-voxel counter %d != oracle %d" % (st["updated_voxels"], ref.total_updated())
    assert int(st["touched_units"]) == ref.total_touched(), "(unit, frame) pairs %d != oracle %d" % (st["touched_units"], ref.total_touched())
    return {"units": int(len(rk)), "tsdf_weight_bit_exact": True, "max_color_diff_0_255": max_dc, "observed_voxels": nz,
            "updated_voxels": int(st["updated_voxels"]), "touched_unit_frames": int(st["touched_units"]), "counters_equal": True}


def mesh_signature(vtx: np.ndarray, tri: np.ndarray, ids: np.ndarray):
    """Canonical relabelling: vertices sorted by global edge id (X, Y, Z, axis); triangles rewritten, rotated so the
    smallest index comes first (orientation kept) and sorted."""
    order = np.lexsort((ids[:, 3], ids[:, 2], ids[:, 1], ids[:, 0]))
    rank = np.empty(len(order), np.int64)
    rank[order] = np.arange(len(order))
    t = rank[tri]
    r = t.argmin(1)
    t = np.stack([t[np.arange(len(t)), (r + k) % 3] for k in range(3)], 1)
    t = t[np.lexsort((t[:, 2], t[:, 1], t[:, 0]))]
    return ids[order], vtx[order], t, order


def mesh_vs_oracle(vol, ref, voxel_length: float) -> dict:
    """extract_triangle_mesh of both: vertex-id set and triangle connectivity + winding bit-exact, positions within
    1e-4 voxel, colours within 1e-5."""
    mesh = vol.extract_triangle_mesh(with_ids=True)
    vtx, vcol, tri, ids = ref.extract_mesh()
    g = mesh_signature(np.asarray(mesh.vertices), np.asarray(mesh.triangles).astype(np.int64), mesh.edge_ids.cpu().numpy())
    r = mesh_signature(vtx, tri.astype(np.int64), ids)
    assert len(g[0]) == len(r[0]) and len(g[2]) == len(r[2]), "mesh size differs: V %d/%d F %d/%d" % (len(g[0]), len(r[0]), len(g[2]), len(r[2]))
    assert np.array_equal(g[0], r[0]), "vertex (edge id) sets differ"
    assert np.array_equal(g[2], r[2]), "triangle connectivity / winding differs"
    dv = float(np.abs(g[1] - r[1]).max()) if len(g[1]) else 0.0
    assert dv <= 1e-4 * voxel_length, "vertex positions differ by %g voxel" % (dv / voxel_length)
    dc = 0.0
    if vcol is not None and len(vcol):
        dc = float(np.abs(np.asarray(mesh.vertex_colors)[g[3]] - vcol[r[3]]).max())
        assert dc < 1e-5, "vertex colours differ by %g" % dc
    return {"vertices": int(len(g[0])), "triangles": int(len(g[2])), "topology_bit_exact": True, "max_vertex_diff_voxel": dv / voxel_length,
            "max_color_diff": dc}


def points_vs_oracle(vol, ref, voxel_length: float) -> dict:
    """extract_point_cloud of both: (voxel, axis) id set bit-exact, positions within 1e-4 voxel, normals 1e-9."""
    pcd = vol.extract_point_cloud(with_ids=True)
    xyz, rgb, nrm, ids = ref.extract_points()
    gi = pcd.edge_ids.cpu().numpy()
    go = np.lexsort((gi[:, 3], gi[:, 2], gi[:, 1], gi[:, 0]))
    ro = np.lexsort((ids[:, 3], ids[:, 2], ids[:, 1], ids[:, 0]))
    assert len(go) == len(ro) and np.array_equal(gi[go], ids[ro]), "surface point (voxel, axis) sets differ"
    dp = float(np.abs(np.asarray(pcd.points)[go] - xyz[ro]).max()) if len(go) else 0.0
    assert dp <= 1e-4 * voxel_length
    fin = np.isfinite(nrm[ro]).all(1)
    dn = float(np.abs(np.asarray(pcd.normals)[go][fin] - nrm[ro][fin]).max()) if fin.any() else 0.0
    assert dn < 1e-9, "normals differ by %g" % dn
    return {"points": int(len(go)), "ids_bit_exact": True, "max_point_diff_voxel": dp / voxel_length, "max_normal_diff": dn}
