"""The allocation-free entry point b2r_mesh_vertex_normals_noalloc (it has no
counterpart in the reference): called through the C-ABI and checked against the existing b2r_mesh_vertex_normals on two calls in
a row (the second mesh is smaller, so stale scratch contents of the first call would show), plus the size limit.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from azurekinect3dreconstruction_b200 import _lib

MAX_V, MAX_F = 262144, 524288          # B2R_NOALLOC_MAX_VERTICES / _TRIANGLES in include/b2recon.h


def grid_mesh(nx, ny, seed):
    """A bumpy height-field mesh: (nx * ny) vertices, 2 (nx - 1)(ny - 1) triangles in a shuffled order, plus one isolated vertex."""
    rng = np.random.default_rng(seed)
    x, y = np.meshgrid(np.arange(nx, dtype=np.float64), np.arange(ny, dtype=np.float64), indexing="ij")
    z = rng.normal(0.0, 0.3, size=x.shape)
    vtx = np.stack([x, y, z], -1).reshape(-1, 3) * 0.004
    vtx = np.concatenate([vtx, [[9.0, 9.0, 9.0]]])                       # belongs to no triangle: its normal stays zero
    i = (np.arange(nx - 1)[:, None] * ny + np.arange(ny - 1)[None, :]).reshape(-1)
    tri = np.concatenate([np.stack([i, i + ny, i + 1], 1), np.stack([i + 1, i + ny, i + ny + 1], 1)]).astype(np.int32)
    rng.shuffle(tri)
    return vtx, tri


def both(lib, vtx, tri):
    dev = torch.device("cuda", torch.cuda.current_device())
    v = torch.as_tensor(vtx, dtype=torch.float64, device=dev).contiguous()
    t = torch.as_tensor(tri, dtype=torch.int32, device=dev).contiguous()
    ref = torch.full((v.shape[0], 3), 7.0, dtype=torch.float64, device=dev)
    out = torch.full((v.shape[0], 3), 7.0, dtype=torch.float64, device=dev)
    s = _lib.stream_ptr()
    _lib.check(lib.b2r_mesh_vertex_normals(_lib.ptr(v), v.shape[0], _lib.ptr(t), t.shape[0], _lib.ptr(ref), s), "vertex_normals")
    _lib.check(lib.b2r_mesh_vertex_normals_noalloc(_lib.ptr(v), v.shape[0], _lib.ptr(t), t.shape[0], _lib.ptr(out), s), "vertex_normals_noalloc")
    torch.cuda.synchronize()
    return ref.cpu().numpy(), out.cpu().numpy()


def test_no_alloc_vertex_normals_match_existing_entry_point_twice():
    lib = _lib.load()
    _lib.require_cuda()
    for nx, ny, seed in ((300, 200, 1), (37, 53, 2)):                    # 60 001 vertices / 119 002 triangles, then a smaller mesh
        vtx, tri = grid_mesh(nx, ny, seed)
        ref, out = both(lib, vtx, tri)
        assert np.isfinite(out).all()
        assert np.abs(out - ref).max() < 1e-9                             # the tolerance the other vertex-normal tests use
        assert np.array_equal(out, ref)                                   # in fact the same sums in the same order
        assert np.all(out[-1] == 0.0) and np.abs(np.linalg.norm(out[:-1], axis=1) - 1.0).max() < 1e-12
    # beyond the fixed scratch: a clear error, nothing launched (the pointers are never dereferenced)
    dev = torch.device("cuda", torch.cuda.current_device())
    dummy = torch.zeros((4, 3), dtype=torch.float64, device=dev)
    tri4 = torch.zeros((1, 3), dtype=torch.int32, device=dev)
    for nV, nF in ((MAX_V + 1, 1), (4, MAX_F + 1)):
        rc = lib.b2r_mesh_vertex_normals_noalloc(_lib.ptr(dummy), nV, _lib.ptr(tri4), nF, _lib.ptr(dummy), _lib.stream_ptr())
        assert rc == _lib.ERR_CAPACITY and "exceeds the fixed scratch" in _lib.last_error()
    # empty mesh is a no-op
    assert lib.b2r_mesh_vertex_normals_noalloc(_lib.ptr(dummy), 0, _lib.ptr(tri4), 0, _lib.ptr(dummy), _lib.stream_ptr()) == 0
