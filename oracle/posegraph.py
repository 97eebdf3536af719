"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy, dense) of Open3D 0.19's pose-graph global optimisation, the
algorithm behind o3d.pipelines.registration.global_optimization as the reference calls it
(/root/reference/CodeThatKindaWorks/capture_frames.py:74-89; CodeThatWorks/bundlefusionattemps/bundlefusion_reconstruction.py:260-273,292-304).

Open3D (pinned 0.19.0 by the reference's environment) is not vendored in /root/reference and cannot be installed
here, so this follows the published algorithm -- pipelines/registration/GlobalOptimization.cpp, function by function, with the
dense H the upstream code builds (Eigen::MatrixXd) and a dense solve.  PARITY UNPINNED: no upstream golden vectors exist for
this path; the anchors are the analytic properties in tests/test_posegraph.py (exact recovery of a consistent graph, rejection
of a false loop closure, gauge fixed on the reference node).

Only tests/ may import this module.  Dictionaries stand in for the graph: nodes = [4x4], edges = [dict(s, t, T, info,
uncertain, conf)].
"""
import copy

import numpy as np


def _op(rows):
    m = np.zeros((4, 4))
    for r, c, v in rows:
        m[r, c] = v
    return m


# jacobian_operator (GlobalOptimization.cpp): generators of alpha, beta, gamma, a, b, c
JACOBIAN_OPERATOR = [_op([(1, 2, -1), (2, 1, 1)]), _op([(0, 2, 1), (2, 0, -1)]), _op([(0, 1, -1), (1, 0, 1)]),
                     _op([(0, 3, 1)]), _op([(1, 3, 1)]), _op([(2, 3, 1)])]


def transform_vector6d_to_matrix4d(v):          # utility/Eigen.cpp
    def rx(a): return np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])
    def ry(a): return np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])
    def rz(a): return np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    m = np.identity(4)
    m[:3, :3] = rz(v[2]) @ ry(v[1]) @ rx(v[0])
    m[:3, 3] = v[3:6]
    return m


def transform_matrix4d_to_vector6d(m):
    r = m[:3, :3]
    sy = np.sqrt(r[0, 0] * r[0, 0] + r[1, 0] * r[1, 0])
    if not (sy < 1e-6):
        x, y, z = np.arctan2(r[2, 1], r[2, 2]), np.arctan2(-r[2, 0], sy), np.arctan2(r[1, 0], r[0, 0])
    else:
        x, y, z = np.arctan2(-r[1, 2], r[1, 1]), np.arctan2(-r[2, 0], sy), 0.0
    return np.array([x, y, z, m[0, 3], m[1, 3], m[2, 3]])


def get_linearized_6d_vector(m):
    return np.array([(-m[1, 2] + m[2, 1]) / 2.0, (-m[2, 0] + m[0, 2]) / 2.0, (-m[0, 1] + m[1, 0]) / 2.0, m[0, 3], m[1, 3], m[2, 3]])


def get_relative_poses(nodes, e):
    return np.linalg.inv(e["T"]), nodes[e["s"]], np.linalg.inv(nodes[e["t"]])


def compute_zeta(nodes, edges):
    out = np.zeros(6 * len(edges))
    for k, e in enumerate(edges):
        x_inv, ts, tt_inv = get_relative_poses(nodes, e)
        out[6 * k:6 * k + 6] = get_linearized_6d_vector(x_inv @ tt_inv @ ts)
    return out


def compute_residual(edges, zeta, lpw):
    r = 0.0
    for k, e in enumerate(edges):
        z = zeta[6 * k:6 * k + 6]
        r += e["conf"] * (z @ e["info"] @ z) + lpw * (np.sqrt(e["conf"]) - 1.0) ** 2
    return r


def update_confidence(edges, zeta, lpw):
    for k, e in enumerate(edges):
        if e["uncertain"]:
            z = zeta[6 * k:6 * k + 6]
            t = lpw / (lpw + z @ e["info"] @ z)
            e["conf"] = t * t


def compute_linear_system(nodes, edges, zeta):
    n = len(nodes)
    H = np.zeros((6 * n, 6 * n))
    b = np.zeros(6 * n)
    for k, e in enumerate(edges):
        z = zeta[6 * k:6 * k + 6]
        x_inv, ts, tt_inv = get_relative_poses(nodes, e)
        js = np.zeros((6, 6))
        jt = np.zeros((6, 6))
        for i in range(6):
            js[:, i] = get_linearized_6d_vector(x_inv @ tt_inv @ JACOBIAN_OPERATOR[i] @ ts)
            jt[:, i] = get_linearized_6d_vector(x_inv @ tt_inv @ -JACOBIAN_OPERATOR[i] @ ts)
        jst_info, jtt_info, et_info, l = js.T @ e["info"], jt.T @ e["info"], z @ e["info"], e["conf"]
        i0, j0 = 6 * e["s"], 6 * e["t"]
        H[i0:i0 + 6, i0:i0 + 6] += l * jst_info @ js
        H[i0:i0 + 6, j0:j0 + 6] += l * jst_info @ jt
        H[j0:j0 + 6, i0:i0 + 6] += l * jtt_info @ js
        H[j0:j0 + 6, j0:j0 + 6] += l * jtt_info @ jt
        b[i0:i0 + 6] -= l * et_info @ js
        b[j0:j0 + 6] -= l * et_info @ jt
    return H, b


def compute_line_process_weight(edges, option):
    if not edges:
        return 0.0
    avg = sum(e["info"][5, 5] for e in edges) / float(len(edges))
    return option["preference_loop_closure"] * option["max_correspondence_distance"] ** 2 * avg


def update_pose_vector(nodes):
    return np.concatenate([transform_matrix4d_to_vector6d(p) for p in nodes]) if nodes else np.zeros(0)


DEFAULT_CRITERIA = dict(max_iteration=100, min_relative_increment=1e-6, min_relative_residual_increment=1e-6, min_right_term=1e-6,
                        min_residual=1e-6, max_iteration_lm=20, upper_scale_factor=2.0 / 3.0, lower_scale_factor=1.0 / 3.0)


def optimize_lm(nodes, edges, criteria, option):
    """GlobalOptimizationLevenbergMarquardt::OptimizePoseGraph; nodes / edges are updated in place. Returns (iterations, lm steps)."""
    n = len(nodes)
    lpw = compute_line_process_weight(edges, option)
    zeta = compute_zeta(nodes, edges)
    current = compute_residual(edges, zeta, lpw)
    update_confidence(edges, zeta, lpw)
    x = update_pose_vector(nodes)
    H, b = compute_linear_system(nodes, edges, zeta)
    lam, ni, rho = 1e-5 * H.diagonal().max(), 2.0, 0.0
    stop = b.max() < criteria["min_right_term"]
    iters = steps = 0
    if stop:
        return iters, steps
    it = 0
    while not stop:
        lm_count = 0
        while True:
            delta = np.linalg.solve(H + lam * np.identity(6 * n), b)
            stop = stop or np.linalg.norm(delta) < criteria["min_relative_increment"] * (np.linalg.norm(x) + criteria["min_relative_increment"])
            if not stop:
                new_nodes = [transform_vector6d_to_matrix4d(delta[6 * i:6 * i + 6]) @ nodes[i] for i in range(n)]
                zeta_new = compute_zeta(new_nodes, edges)
                new_res = compute_residual(edges, zeta_new, lpw)
                rho = (current - new_res) / (delta @ (lam * delta + b) + 1e-3)
                if rho > 0:
                    stop = stop or (current - new_res < criteria["min_relative_residual_increment"] * current)
                    if stop:
                        break
                    alpha = min(1.0 - (2 * rho - 1) ** 3, criteria["upper_scale_factor"])
                    lam *= max(criteria["lower_scale_factor"], alpha)
                    ni = 2.0
                    current = new_res
                    zeta = zeta_new
                    nodes[:] = new_nodes
                    x = update_pose_vector(nodes)
                    update_confidence(edges, zeta, lpw)
                    H, b = compute_linear_system(nodes, edges, zeta)
                    stop = stop or b.max() < criteria["min_right_term"]
                    if stop:
                        break
                else:
                    lam *= ni
                    ni *= 2
            lm_count += 1
            steps += 1
            stop = stop or lm_count >= criteria["max_iteration_lm"]
            if rho > 0 or stop:
                break
        iters += 1
        stop = stop or current < criteria["min_residual"] or it >= criteria["max_iteration"]
        it += 1
    return iters, steps


def global_optimization(nodes, edges, option, criteria=None):
    """GlobalOptimization(): returns (nodes, edges) of the optimised, pruned graph; inputs are not modified."""
    criteria = dict(DEFAULT_CRITERIA if criteria is None else criteria)
    pre_nodes = [np.array(p, dtype=np.float64) for p in nodes]
    pre_edges = copy.deepcopy(edges)
    r1 = optimize_lm(pre_nodes, pre_edges, criteria, option)
    pruned = [e for e in pre_edges if (not e["uncertain"]) or e["conf"] > option["edge_prune_threshold"]]
    r2 = optimize_lm(pre_nodes, pruned, criteria, option)
    ref = option.get("reference_node", -1)
    if 0 <= ref < len(nodes):
        comp = np.asarray(nodes[ref]) @ np.linalg.inv(pre_nodes[ref])
        pre_nodes = [comp @ p for p in pre_nodes]
    return pre_nodes, pruned, (r1, r2)
