"""Host-side logic of the multi-GPU path on CPU: world_size-2 gloo process group.
The kernels are exercised on the GPU in test_gpu_parity.py::test_sharded_extraction_equals_single."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from azurekinect3dreconstruction_b200 import _lib
from azurekinect3dreconstruction_b200.pipelines import sharded


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = sharded.Comm()
        assert (comm.rank, comm.world) == (rank, world)
        # ragged gather (rank 0: 3 rows, rank 1: 0 rows) of a 2-D int tensor and a 1-D float tensor
        a = torch.arange(6, dtype=torch.int32).reshape(3, 2) + 100 * rank if rank == 0 else torch.zeros((0, 2), dtype=torch.int32)
        got = comm.all_gather_v(a)
        assert [g.shape[0] for g in got] == [3, 0] and torch.equal(got[0], torch.arange(6, dtype=torch.int32).reshape(3, 2))
        b = torch.full((rank + 2,), float(rank), dtype=torch.float64)
        got = comm.all_gather_v(b)
        assert [g.shape[0] for g in got] == [2, 3] and float(got[1].sum()) == 3.0
        # mesh merge: each rank contributes vertices [base, base+n) in rank-major global numbering
        n = 4 + rank
        base = 0 if rank == 0 else 4
        vtx = torch.arange(3 * n, dtype=torch.float64).reshape(n, 3) + 1000 * rank
        tri = torch.tensor([[base, base + 1, (base + n) % 9]], dtype=torch.int32)      # may reference the other rank
        parts = [comm.all_gather_v(vtx), comm.all_gather_v(tri)]
        mesh = sharded.merge_mesh_parts(parts)
        V, F = np.asarray(mesh.vertices), np.asarray(mesh.triangles)
        assert V.shape == (9, 3) and F.shape == (2, 3) and V[4, 0] == 1000.0 and F[1, 0] == 4
        # every unit key has exactly one owner; the two ranks partition a key block evenly enough
        keys = [(x, y, z) for x in range(-6, 6) for y in range(-3, 3) for z in range(-6, 6)]
        mine = sum(1 for k in keys if sharded.unit_owner(k, world) == rank)
        cnt = torch.tensor([mine], dtype=torch.int64)
        dist.all_reduce(cnt)
        assert int(cnt.item()) == len(keys) and 0.35 * len(keys) < mine < 0.65 * len(keys)
        # ragged all-to-all and gather-v (slab re-partition of the extraction): rank r sends r + d + 1 rows to rank d
        parts = [torch.full((rank + d + 1, 2), 10 * rank + d, dtype=torch.float32) for d in range(world)]
        got = comm.all_to_all_v(parts)
        assert [g.shape[0] for g in got] == [src + rank + 1 for src in range(world)]
        assert all(float(g[0, 0]) == 10 * src + rank for src, g in enumerate(got))
        gv = comm.gather_v(torch.full((rank + 1,), float(rank)), 0)
        assert (gv is None) if rank else ([g.shape[0] for g in gv] == [1, 2] and float(gv[1][0]) == 1.0)
        # host-frame exchange: each rank uploads its share, the all-gather returns every frame in global order
        F, H, W = 11, 4, 6                                                       # ragged: 11 frames, super-chunks of 2 x 3
        depth = (torch.arange(F * H * W, dtype=torch.int32) % 60000).to(torch.int16).reshape(F, H, W).view(torch.uint16)
        color = (torch.arange(F * H * W * 3, dtype=torch.int64) % 251).to(torch.uint8).reshape(F, H, W, 3)
        import ctypes as C
        for ramp, expect in ((False, [(0, 6), (6, 5)]), (True, [(0, 2), (2, 4), (6, 5)])):
            ex = sharded.make_frame_exchange((H, W), 3, "nccl", per=3, device=None, ramp=ramp, mode="pull")   # the collective form ("nccl" = torch.distributed: gloo here)
            assert isinstance(ex, sharded.HostFrameExchange)
            seen = []

            def consume(dp, cp, f0, n):
                assert len(dp) == len(cp) == n
                d = [np.frombuffer((C.c_uint8 * (H * W * 2)).from_address(p), dtype=np.int16).copy().reshape(H, W) for p in dp]
                c = [np.frombuffer((C.c_uint8 * (H * W * 3)).from_address(p), dtype=np.uint8).copy().reshape(H, W, 3) for p in cp]
                seen.append((f0, n, np.stack(d), np.stack(c)))
            ex.run(depth, color, consume)
            assert [(f0, n) for f0, n, _, _ in seen] == expect
            got_d = np.concatenate([d for _, _, d, _ in seen])
            got_c = np.concatenate([c for _, _, _, c in seen])
            assert np.array_equal(got_d, depth.view(torch.int16).numpy()) and np.array_equal(got_c, color.numpy())
            plan = sharded.superchunk_plan(F, world, 3, ramp)
            assert ex.bytes_uploaded == sum(f1 - f0 for _, _, rg in plan for f0, f1 in [rg[rank]]) * H * W * 5
        # config 5's stream exchange: 4 streams (2 per rank), 5 ticks in batches of 2 -> every rank sees every frame, tick-major
        S, ticks, H5, W5 = 4, 5, 3, 8
        tex = sharded.StreamTickExchange((H5, W5), 3, S, ticks_per_batch=2, device=None)
        assert (tex.rank, tex.world, tex.s_local) == (rank, world, 2)
        fd = lambda t, s: torch.full((H5, W5), 1000 * t + s, dtype=torch.int16).view(torch.uint16)
        fc = lambda t, s: torch.full((H5, W5, 3), (7 * t + s) % 256, dtype=torch.uint8)
        fe = lambda t, s: np.eye(4) * (1 + t) + s
        mine = [s for s in range(S) if s // tex.s_local == rank]
        packed = torch.stack([tex.pack(torch.stack([fd(t, s) for s in mine]), torch.stack([fc(t, s) for s in mine]), np.stack([fe(t, s) for s in mine]))
                              for t in range(ticks)])
        assert packed.shape == (ticks, 2, tex.rec) and tex.off_color % 256 == 0 and tex.off_pose % 256 == 0
        got = []

        def consume(dp, cp, E, t0, t):
            assert len(dp) == len(cp) == t * S and E.shape == (t * S, 4, 4)
            for i in range(t * S):
                d = np.frombuffer((C.c_uint8 * (H5 * W5 * 2)).from_address(dp[i]), dtype=np.uint16).copy()
                c = np.frombuffer((C.c_uint8 * (H5 * W5 * 3)).from_address(cp[i]), dtype=np.uint8).copy()
                got.append((t0 + i // S, i % S, int(d[0]), int(d[-1]), int(c[0]), E[i].copy()))
        tex.run(packed, consume)
        assert [(t, s) for t, s, *_ in got] == [(t, s) for t in range(ticks) for s in range(S)]
        for t, s, d0, d1, c0, E in got:
            assert d0 == d1 == 1000 * t + s and c0 == (7 * t + s) % 256 and np.array_equal(E, fe(t, s))
        assert tex.bytes_gathered == ticks * S * tex.rec
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_gather_merge_and_ownership():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == {0: "ok", 1: "ok"}, res


def test_unit_owner_python_mirror_matches_library():
    lib = _lib.load()
    rng = np.random.default_rng(5)
    for world in (1, 2, 3, 4, 8):
        for k in rng.integers(-5000, 5000, size=(200, 3)):
            assert lib.b2r_unit_owner(int(k[0]), int(k[1]), int(k[2]), world) == sharded.unit_owner(k, world)
    assert lib.b2r_unit_owner(1 << 21, 0, 0, 4) == -1


def test_superchunk_plan_keeps_global_frame_order():
    for F, world, per, ramp in ((300, 8, 8, False), (11, 2, 3, False), (5, 4, 2, False), (64, 8, 8, False), (1, 2, 4, False),
                                (300, 8, 8, True), (11, 2, 3, True), (7, 4, 4, True), (300, 2, 16, True)):
        plan = sharded.superchunk_plan(F, world, per, ramp)
        covered = []
        for f_start, n_valid, ranges in plan:
            assert len(ranges) == world and 0 < n_valid <= world * per
            slot = 0
            for r, (f0, f1) in enumerate(ranges):
                assert 0 <= f1 - f0 <= per
                if f1 > f0:
                    cur = ranges[0][1] - ranges[0][0]
                    assert f0 - f_start == r * cur                              # row in the gathered buffer == frame index
                covered += list(range(f0, f1))
            assert covered[-n_valid:] == list(range(f_start, f_start + n_valid))
        assert covered == list(range(F))


def test_slab_cuts_and_routes():
    rng = np.random.default_rng(3)
    for world in (1, 2, 3, 4, 8):
        kx = rng.integers(-7, 9, size=500).astype(np.int32)
        slabs = sharded.slab_cuts(kx, world)
        assert len(slabs) == world
        covered = np.zeros(kx.shape, np.int32)
        prev_hi = None
        for lo, hi in slabs:
            if hi < lo:
                continue
            assert prev_hi is None or lo > prev_hi                            # disjoint, ascending
            prev_hi = hi
            covered += ((kx >= lo) & (kx <= hi))
        assert np.all(covered == 1)                                           # every unit layer belongs to exactly one slab
        sizes = [int(((kx >= lo) & (kx <= hi)).sum()) for lo, hi in slabs if hi >= lo]
        assert max(sizes) <= 500 / min(world, 16) + 80                        # balanced up to one layer of units
        routes = sharded.slab_routes(kx, slabs)
        for (lo, hi), ids in zip(slabs, routes):
            if hi < lo:
                assert len(ids) == 0
            else:
                assert set(np.unique(kx[ids]).tolist()) == set(v for v in np.unique(kx).tolist() if lo - 1 <= v <= hi + 2)
    assert sharded.slab_cuts(np.zeros((0,), np.int32), 3) == [(1, 0)] * 3
