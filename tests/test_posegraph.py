"""f3: PoseGraph + global_optimization (host LM in libb2recon.so) against the numpy restatement in oracle/posegraph.py
and against analytic properties; pose-graph JSON layout.  CPU tests: the solver is host code by design."""
import json
import os
import sys
import time
import warnings

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import posegraph as ref                                                   # noqa: E402  (the checker)

import azurekinect3dreconstruction_b200 as o3d                             # noqa: E402
from azurekinect3dreconstruction_b200 import synthetic as syn              # noqa: E402

R = o3d.pipelines.registration


def small_motion(rng, rot=0.01, trans=0.01):
    return ref.transform_vector6d_to_matrix4d(np.concatenate([rng.normal(0, rot, 3), rng.normal(0, trans, 3)]))


def make_graph(n=40, seed=0, odo_noise=(0.002, 0.003), closures=((0, 39), (5, 30)), false_closure=None, info_scale=2000.0):
    """Ground-truth orbit; odometry edges carry noisy relative transforms, node poses are the chained (drifting) estimates;
    loop closures carry the TRUE relative transform and are uncertain.  Edge (s, t, T) with T = Tt^-1 Ts (upstream's residual)."""
    rng = np.random.default_rng(seed)
    gt = syn.orbit_trajectory(n)
    info = np.identity(6) * info_scale
    pg = R.PoseGraph()
    est = [gt[0].copy()]
    pg.nodes.append(R.PoseGraphNode(est[0]))
    for i in range(1, n):
        T = np.linalg.inv(gt[i]) @ gt[i - 1] @ small_motion(rng, *odo_noise)
        est.append(est[i - 1] @ np.linalg.inv(T))                         # Ts = Tt T  ->  Tt = Ts T^-1
        pg.nodes.append(R.PoseGraphNode(est[i]))
        pg.edges.append(R.PoseGraphEdge(i - 1, i, T, info, uncertain=False))
    for s, t in closures:
        pg.edges.append(R.PoseGraphEdge(s, t, np.linalg.inv(gt[t]) @ gt[s], info, uncertain=True))
    if false_closure:
        s, t = false_closure
        wrong = np.linalg.inv(gt[t]) @ gt[s] @ ref.transform_vector6d_to_matrix4d(np.array([0.3, -0.2, 0.4, 0.5, -0.4, 0.3]))
        pg.edges.append(R.PoseGraphEdge(s, t, wrong, info, uncertain=True))
    return pg, gt


def as_ref(pg):
    nodes = [nd.pose.copy() for nd in pg.nodes]
    edges = [dict(s=e.source_node_id, t=e.target_node_id, T=e.transformation.copy(), info=e.information.copy(), uncertain=e.uncertain, conf=e.confidence)
             for e in pg.edges]
    return nodes, edges


def option(**kw):
    d = dict(max_correspondence_distance=0.05, edge_prune_threshold=0.25, preference_loop_closure=2.0, reference_node=0)
    d.update(kw)
    return d


def pose_error(pg, gt):
    return max(np.abs(nd.pose - g).max() for nd, g in zip(pg.nodes, gt))


def test_consistent_graph_is_recovered_exactly_and_gauge_stays_on_reference_node():
    pg, gt = make_graph(odo_noise=(0.0, 0.0))
    rng = np.random.default_rng(3)
    for nd in pg.nodes[1:]:
        nd.pose = small_motion(rng, 0.02, 0.03) @ nd.pose                  # disturb the estimates; every edge is exact
    assert pose_error(pg, gt) > 1e-2
    R.global_optimization(pg, R.GlobalOptimizationLevenbergMarquardt(), R.GlobalOptimizationConvergenceCriteria(), R.GlobalOptimizationOption(**option()))
    assert pg.last_report["valid"] and len(pg.edges) == 41
    assert np.array_equal(pg.nodes[0].pose, gt[0]) or np.abs(pg.nodes[0].pose - gt[0]).max() < 1e-12
    assert pose_error(pg, gt) < 1e-4                                       # the linearised SE(3) residual stops within its criteria


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_matches_numpy_restatement_on_noisy_graph_with_loop_closures(seed):
    pg, gt = make_graph(seed=seed, false_closure=(3, 25) if seed else None)
    nodes, edges = as_ref(pg)
    before = pose_error(pg, gt)
    want_nodes, want_edges, (r1, r2) = ref.global_optimization(nodes, edges, option())
    R.global_optimization(pg, R.GlobalOptimizationLevenbergMarquardt(), R.GlobalOptimizationConvergenceCriteria(), R.GlobalOptimizationOption(**option()))
    rep = pg.last_report
    assert rep["iterations"] == (r1[0], r2[0]) and rep["lm_steps"] == (r1[1], r2[1])
    assert [(e.source_node_id, e.target_node_id) for e in pg.edges] == [(e["s"], e["t"]) for e in want_edges]
    for nd, w in zip(pg.nodes, want_nodes):
        np.testing.assert_allclose(nd.pose, w, rtol=0, atol=1e-9)          # envelope Cholesky vs dense LU: same systems, different elimination order
    np.testing.assert_allclose([e.confidence for e in pg.edges], [e["conf"] for e in want_edges], rtol=0, atol=1e-9)
    assert rep["residual_after"] < rep["residual_before"]
    assert pose_error(pg, gt) < before                                     # closing the loop removes drift
    if seed:
        assert len(pg.edges) == len(edges) - 1                             # the false closure fell under the prune threshold


def test_invalid_graphs_are_left_untouched_with_a_warning():
    pg, _ = make_graph(n=10, closures=())
    pg.edges.pop(4)                                                        # two components
    poses = [nd.pose.copy() for nd in pg.nodes]
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        R.global_optimization(pg, R.GlobalOptimizationLevenbergMarquardt(), R.GlobalOptimizationConvergenceCriteria(), R.GlobalOptimizationOption())
    assert w and "not connected" in str(w[0].message) and not pg.last_report["valid"]
    assert all(np.array_equal(a, nd.pose) for a, nd in zip(poses, pg.nodes)) and len(pg.edges) == 8
    pg, _ = make_graph(n=10, closures=())
    pg.edges[2].confidence = 0.5                                           # a certain edge must carry confidence 1
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        R.global_optimization(pg, R.GlobalOptimizationLevenbergMarquardt(), R.GlobalOptimizationConvergenceCriteria(), R.GlobalOptimizationOption())
    assert w and "certain edge" in str(w[0].message)
    with pytest.raises(NotImplementedError):
        R.global_optimization(pg, R.GlobalOptimizationGaussNewton(), R.GlobalOptimizationConvergenceCriteria(), R.GlobalOptimizationOption())


def test_long_chain_with_closures_uses_the_envelope_not_a_dense_matrix():
    """2000 nodes (12000 unknowns): a dense factorisation is ~6e11 flops per LM step; the block envelope takes milliseconds."""
    pg, gt = make_graph(n=2000, seed=5, odo_noise=(0.0005, 0.0005), closures=tuple((i, i + 150) for i in range(0, 1800, 100)))
    before = pose_error(pg, gt)
    t0 = time.perf_counter()
    R.global_optimization(pg, R.GlobalOptimizationLevenbergMarquardt(), R.GlobalOptimizationConvergenceCriteria(), R.GlobalOptimizationOption(**option()))
    dt = time.perf_counter() - t0
    assert pg.last_report["valid"] and dt < 20.0
    assert pg.last_report["residual_after"] < pg.last_report["residual_before"]
    assert pose_error(pg, gt) < before


def test_pose_graph_json_layout_and_round_trip(tmp_path):
    pg, _ = make_graph(n=6, closures=((0, 5),))
    pg.edges[-1].confidence = 0.75
    path = str(tmp_path / "pose_graph.json")
    assert o3d.io.write_pose_graph(path, pg)
    doc = json.load(open(path))
    assert doc["class_name"] == "PoseGraph" and doc["version_major"] == 1 and doc["version_minor"] == 0
    n0 = doc["nodes"][3]
    assert n0["class_name"] == "PoseGraphNode" and n0["pose"][12:15] == [float(v) for v in pg.nodes[3].pose[:3, 3]]   # Eigen column-major
    e = doc["edges"][-1]
    assert e["class_name"] == "PoseGraphEdge" and e["uncertain"] is True and e["confidence"] == 0.75 and len(e["information"]) == 36
    assert (e["source_node_id"], e["target_node_id"]) == (0, 5)
    back = o3d.io.read_pose_graph(path)
    assert len(back.nodes) == 6 and len(back.edges) == 6
    for a, b in zip(pg.nodes, back.nodes):
        assert np.array_equal(a.pose, b.pose)
    for a, b in zip(pg.edges, back.edges):
        assert np.array_equal(a.transformation, b.transformation) and np.array_equal(a.information, b.information)
        assert (a.source_node_id, a.target_node_id, a.uncertain, a.confidence) == (b.source_node_id, b.target_node_id, b.uncertain, b.confidence)
