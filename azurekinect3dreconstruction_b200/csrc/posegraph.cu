// posegraph.cu -- SURVEY 8(f) row f3: PoseGraph global optimisation, the consumer of a12's information matrix.
//
// Replaces, behind the C-ABI in include/b2recon.h, what the reference reaches through
//   o3d.pipelines.registration.global_optimization(pose_graph, GlobalOptimizationLevenbergMarquardt(),
//                                                  GlobalOptimizationConvergenceCriteria(), GlobalOptimizationOption(...))
// (/root/reference/CodeThatKindaWorks/capture_frames.py:74-89; CodeThatWorks/bundlefusionattemps/
//  bundlefusion_reconstruction.py:260-273,292-304).  Semantics: Open3D 0.19 pipelines/registration/GlobalOptimization.cpp
// (Choi et al. 2015 line processes): validate -> LM on a copy -> prune uncertain edges below the threshold -> LM again ->
// re-anchor on the reference node.
//
// HOST code by design ("small sparse LM on host"): 6 N unknowns with N a few hundred to a few thousand
// and tens of dependent LM steps of ~100 us each -- a kernel launch per step would cost more than the step.  What is
// B200-shaped here is the solver layout: the normal matrix is kept as a block envelope (skyline) under a reverse
// Cuthill-McKee ordering of the nodes, so an odometry chain with loop closures factorises in O(N * band^2) instead of
// the dense O(N^3) of the reference's Eigen::MatrixXd H.  The GPU half of f3 is the re-integration of every frame under
// the optimised poses (volume.reset() + the batched integrate path of tsdf.cu).
#include "common.cuh"
#include "host_math.h"
#include <vector>
#include <algorithm>
#include <queue>

namespace {

struct Edge {
    int s, t;
    double T[16];      // transformation, row-major
    double info[36];   // information, row-major (symmetric in practice; used as given)
    int uncertain;
    double conf;
};
struct Graph {
    std::vector<double> pose;   // n x 16, row-major
    std::vector<Edge> edges;
    int n() const { return (int)(pose.size() / 16); }
};
struct Option { double max_corr, prune, pref_loop; int reference; };
struct Criteria {
    int max_iteration; double min_relative_increment, min_relative_residual_increment, min_right_term, min_residual;
    int max_iteration_lm; double upper_scale_factor, lower_scale_factor;
};

// utility::TransformVector6dToMatrix4d: R = Rz(v2) * Ry(v1) * Rx(v0), t = v[3..5]
void vec6_to_m4(const double* v, double* m) {
    const double cx = std::cos(v[0]), sx = std::sin(v[0]), cy = std::cos(v[1]), sy = std::sin(v[1]), cz = std::cos(v[2]), sz = std::sin(v[2]);
    m[0] = cz * cy; m[1] = cz * sy * sx - sz * cx; m[2] = cz * sy * cx + sz * sx; m[3] = v[3];
    m[4] = sz * cy; m[5] = sz * sy * sx + cz * cx; m[6] = sz * sy * cx - cz * sx; m[7] = v[4];
    m[8] = -sy;     m[9] = cy * sx;                m[10] = cy * cx;               m[11] = v[5];
    m[12] = 0; m[13] = 0; m[14] = 0; m[15] = 1;
}
// utility::TransformMatrix4dToVector6d
void m4_to_vec6(const double* m, double* v) {
    const double sy = std::sqrt(m[0] * m[0] + m[4] * m[4]);
    if (!(sy < 1e-6)) { v[0] = std::atan2(m[9], m[10]); v[1] = std::atan2(-m[8], sy); v[2] = std::atan2(m[4], m[0]); }
    else { v[0] = std::atan2(-m[6], m[5]); v[1] = std::atan2(-m[8], sy); v[2] = 0.0; }
    v[3] = m[3]; v[4] = m[7]; v[5] = m[11];
}
// GetLinearized6DVector
inline void linearized6(const double* m, double* o) {
    o[0] = (-m[6] + m[9]) / 2.0; o[1] = (-m[8] + m[2]) / 2.0; o[2] = (-m[1] + m[4]) / 2.0;
    o[3] = m[3]; o[4] = m[7]; o[5] = m[11];
}
// X_inv * Tt_inv and Ts of an edge (GetRelativePoses)
bool relative(const Graph& g, const Edge& e, double* A /* X_inv * Tt_inv */, const double** Ts) {
    double Xi[16], Tti[16];
    if (!b2r::m4_inverse(e.T, Xi) || !b2r::m4_inverse(&g.pose[16 * (size_t)e.t], Tti)) return false;
    b2r::m4_mul(Xi, Tti, A);
    *Ts = &g.pose[16 * (size_t)e.s];
    return true;
}
// ComputeZeta
bool compute_zeta(const Graph& g, std::vector<double>& zeta) {
    zeta.resize(6 * g.edges.size());
    for (size_t k = 0; k < g.edges.size(); k++) {
        double A[16], M[16]; const double* Ts;
        if (!relative(g, g.edges[k], A, &Ts)) return false;
        b2r::m4_mul(A, Ts, M);
        linearized6(M, &zeta[6 * k]);
    }
    return true;
}
inline double quad6(const double* e, const double* info) {   // e^T * info * e
    double r = 0;
    for (int i = 0; i < 6; i++) { double s = 0; for (int j = 0; j < 6; j++) s += info[6 * i + j] * e[j]; r += e[i] * s; }
    return r;
}
// ComputeResidual (Choi et al. Eq. 9)
double compute_residual(const Graph& g, const std::vector<double>& zeta, double lpw) {
    double r = 0;
    for (size_t k = 0; k < g.edges.size(); k++) {
        const double l = g.edges[k].conf, d = std::sqrt(l) - 1.0;
        r += l * quad6(&zeta[6 * k], g.edges[k].info) + lpw * d * d;
    }
    return r;
}
// UpdateConfidence (Eq. 2)
void update_confidence(Graph& g, const std::vector<double>& zeta, double lpw) {
    for (size_t k = 0; k < g.edges.size(); k++) {
        Edge& e = g.edges[k];
        if (!e.uncertain) continue;
        const double t = lpw / (lpw + quad6(&zeta[6 * k], e.info));
        e.conf = t * t;
    }
}
// ComputeLineProcessWeight (Section 5)
double line_process_weight(const Graph& g, const Option& o) {
    if (g.edges.empty()) return 0.0;
    double avg = 0;
    for (const Edge& e : g.edges) avg += e.info[35];
    avg /= (double)g.edges.size();
    return o.pref_loop * o.max_corr * o.max_corr * avg;
}

// ---- block-envelope normal equations --------------------------------------------------------------------------
// Nodes are renumbered by perm (reverse Cuthill-McKee); scalar row i of the permuted system stores columns
// first[i] .. i contiguously at val[off[i] ...].
struct Envelope {
    int n = 0;
    std::vector<int> perm, first;       // perm[node] = position
    std::vector<size_t> off;
    std::vector<double> val, b;
    double& at(int i, int j) { return val[off[i] + (size_t)(j - first[i])]; }   // j in [first[i], i]
};
void build_layout(const Graph& g, Envelope& E) {
    const int N = g.n();
    std::vector<std::vector<int>> adj(N);
    for (const Edge& e : g.edges) if (e.s != e.t) { adj[e.s].push_back(e.t); adj[e.t].push_back(e.s); }
    for (auto& a : adj) { std::sort(a.begin(), a.end()); a.erase(std::unique(a.begin(), a.end()), a.end()); }
    std::vector<int> order; order.reserve(N);
    std::vector<char> seen(N, 0);
    for (int pass = 0; (int)order.size() < N; pass++) {
        int start = -1;                                  // lowest-degree unvisited node starts a component
        for (int i = 0; i < N; i++) if (!seen[i] && (start < 0 || adj[i].size() < adj[start].size())) start = i;
        std::queue<int> q; q.push(start); seen[start] = 1;
        while (!q.empty()) {
            const int u = q.front(); q.pop(); order.push_back(u);
            std::vector<int> nb;
            for (int v : adj[u]) if (!seen[v]) { seen[v] = 1; nb.push_back(v); }
            std::sort(nb.begin(), nb.end(), [&](int a, int b) { return adj[a].size() != adj[b].size() ? adj[a].size() < adj[b].size() : a < b; });
            for (int v : nb) q.push(v);
        }
    }
    std::reverse(order.begin(), order.end());
    E.perm.assign(N, 0);
    for (int p = 0; p < N; p++) E.perm[order[p]] = p;
    std::vector<int> fb(N);
    for (int i = 0; i < N; i++) fb[i] = i;
    for (const Edge& e : g.edges) {
        const int a = E.perm[e.s], b = E.perm[e.t];
        fb[std::max(a, b)] = std::min(fb[std::max(a, b)], std::min(a, b));
    }
    E.n = 6 * N;
    E.first.resize(E.n); E.off.resize(E.n + 1);
    size_t tot = 0;
    for (int i = 0; i < E.n; i++) { E.first[i] = 6 * fb[i / 6]; E.off[i] = tot; tot += (size_t)(i - E.first[i] + 1); }
    E.off[E.n] = tot;
    E.val.assign(tot, 0.0); E.b.assign(E.n, 0.0);
}
// GetJacobian + ComputeLinearSystem: H += l J^T info J (four 6x6 blocks per edge), b -= l J^T info e
bool compute_system(const Graph& g, const std::vector<double>& zeta, Envelope& E) {
    std::fill(E.val.begin(), E.val.end(), 0.0);
    std::fill(E.b.begin(), E.b.end(), 0.0);
    for (size_t k = 0; k < g.edges.size(); k++) {
        const Edge& e = g.edges[k];
        double A[16]; const double* Ts;
        if (!relative(g, e, A, &Ts)) return false;
        double Js[36], Jt[36];   // row-major 6x6
        for (int c = 0; c < 6; c++) {
            // jacobian_operator[c] * Ts: rotation generators pick/negate rows of Ts, translation generators copy row 3
            double G[16] = {0};
            if (c == 0) { for (int j = 0; j < 4; j++) { G[4 + j] = -Ts[8 + j]; G[8 + j] = Ts[4 + j]; } }
            else if (c == 1) { for (int j = 0; j < 4; j++) { G[j] = Ts[8 + j]; G[8 + j] = -Ts[j]; } }
            else if (c == 2) { for (int j = 0; j < 4; j++) { G[j] = -Ts[4 + j]; G[4 + j] = Ts[j]; } }
            else { for (int j = 0; j < 4; j++) G[4 * (c - 3) + j] = Ts[12 + j]; }
            double M[16], col[6];
            b2r::m4_mul(A, G, M);
            linearized6(M, col);
            for (int r = 0; r < 6; r++) { Js[6 * r + c] = col[r]; Jt[6 * r + c] = -col[r]; }
        }
        const double l = e.conf;
        double JsI[36], JtI[36], eI[6];   // J^T * info, e^T * info
        for (int i = 0; i < 6; i++) {
            for (int j = 0; j < 6; j++) {
                double a = 0, b = 0;
                for (int m = 0; m < 6; m++) { a += Js[6 * m + i] * e.info[6 * m + j]; b += Jt[6 * m + i] * e.info[6 * m + j]; }
                JsI[6 * i + j] = a; JtI[6 * i + j] = b;
            }
            double s = 0;
            for (int m = 0; m < 6; m++) s += zeta[6 * k + m] * e.info[6 * m + i];
            eI[i] = s;
        }
        const int bi = 6 * E.perm[e.s], bj = 6 * E.perm[e.t];
        auto add_block = [&](int r0, int c0, const double* L, const double* R) {   // H[r0.., c0..] += l * L * R (lower triangle kept)
            for (int i = 0; i < 6; i++)
                for (int j = 0; j < 6; j++) {
                    const int r = r0 + i, c = c0 + j;
                    if (c > r) continue;
                    double s = 0;
                    for (int m = 0; m < 6; m++) s += L[6 * i + m] * R[6 * m + j];
                    E.at(r, c) += l * s;
                }
        };
        add_block(bi, bi, JsI, Js);
        add_block(bj, bj, JtI, Jt);
        if (bi > bj) add_block(bi, bj, JsI, Jt); else if (bj > bi) add_block(bj, bi, JtI, Js);
        else { add_block(bi, bi, JsI, Jt); add_block(bi, bi, JtI, Js); }            // self edge: both cross terms land on the diagonal block
        for (int i = 0; i < 6; i++) {
            double s = 0, t = 0;
            for (int m = 0; m < 6; m++) { s += eI[m] * Js[6 * m + i]; t += eI[m] * Jt[6 * m + i]; }
            E.b[bi + i] -= l * s; E.b[bj + i] -= l * t;
        }
    }
    return true;
}
// (H + lambda I) x = b by envelope Cholesky; L overwrites a copy.  false when a pivot is not positive.
bool solve_damped(const Envelope& E, double lambda, std::vector<double>& L, std::vector<double>& x) {
    const int n = E.n;
    L = E.val;
    for (int i = 0; i < n; i++) L[E.off[i] + (size_t)(i - E.first[i])] += lambda;
    for (int i = 0; i < n; i++) {
        double* Li = &L[E.off[i]] - E.first[i];            // Li[j] = L(i, j)
        for (int j = E.first[i]; j <= i; j++) {
            const double* Lj = &L[E.off[j]] - E.first[j];
            double s = Li[j];
            for (int k = std::max(E.first[i], E.first[j]); k < j; k++) s -= Li[k] * Lj[k];
            if (j < i) Li[j] = s / Lj[j];
            else { if (!(s > 0.0)) return false; Li[i] = std::sqrt(s); }
        }
    }
    x = E.b;
    for (int i = 0; i < n; i++) {                           // L y = b
        const double* Li = &L[E.off[i]] - E.first[i];
        double s = x[i];
        for (int k = E.first[i]; k < i; k++) s -= Li[k] * x[k];
        x[i] = s / Li[i];
    }
    for (int i = n - 1; i >= 0; i--) {                      // L^T x = y
        const double* Li = &L[E.off[i]] - E.first[i];
        x[i] /= Li[i];
        for (int k = E.first[i]; k < i; k++) x[k] -= Li[k] * x[i];
    }
    return true;
}

struct Report { int iterations = 0, lm_steps = 0; double residual0 = 0, residual = 0, lambda = 0; };

// GlobalOptimizationLevenbergMarquardt::OptimizePoseGraph
int optimize_lm(Graph& g, const Criteria& c, const Option& o, Report& rep) {
    const int N = g.n();
    const double lpw = line_process_weight(g, o);
    std::vector<double> zeta, zeta_new, delta(6 * (size_t)N), Lbuf, xs(6 * (size_t)N);
    if (!compute_zeta(g, zeta)) return B2R_ERR_SINGULAR;
    double current = compute_residual(g, zeta, lpw), fresh = current;
    rep.residual0 = current;
    update_confidence(g, zeta, lpw);
    Envelope E;
    build_layout(g, E);
    auto pose_vector_norm = [&]() { double s = 0; for (int i = 0; i < N; i++) { double v[6]; m4_to_vec6(&g.pose[16 * (size_t)i], v); for (int k = 0; k < 6; k++) s += v[k] * v[k]; } return std::sqrt(s); };
    auto right_term_small = [&]() { double m = -INFINITY; for (double v : E.b) m = std::max(m, v); return m < c.min_right_term; };
    double xnorm = pose_vector_norm();
    if (!compute_system(g, zeta, E)) return B2R_ERR_SINGULAR;
    double hmax = -INFINITY;
    for (int i = 0; i < E.n; i++) hmax = std::max(hmax, E.val[E.off[i] + (size_t)(i - E.first[i])]);
    double lambda = 1e-5 * hmax, ni = 2.0, rho = 0.0;
    bool stop = right_term_small();
    rep.residual = current; rep.lambda = lambda;
    if (stop) return B2R_OK;
    Graph trial;
    for (int iter = 0; !stop; iter++) {
        int lm_count = 0;
        do {
            std::vector<double> xp;
            if (!solve_damped(E, lambda, Lbuf, xp)) { b2r_set_error("global_optimization: normal matrix not positive definite"); return B2R_ERR_SINGULAR; }
            double dn = 0;
            for (int i = 0; i < N; i++) for (int k = 0; k < 6; k++) { const double v = xp[6 * (size_t)E.perm[i] + k]; delta[6 * (size_t)i + k] = v; dn += v * v; }
            stop = stop || (std::sqrt(dn) < c.min_relative_increment * (xnorm + c.min_relative_increment));
            if (!stop) {
                trial = g;                                                    // UpdatePoseGraph
                for (int i = 0; i < N; i++) { double D[16]; vec6_to_m4(&delta[6 * (size_t)i], D); b2r::m4_mul(D, &g.pose[16 * (size_t)i], &trial.pose[16 * (size_t)i]); }
                if (!compute_zeta(trial, zeta_new)) return B2R_ERR_SINGULAR;
                fresh = compute_residual(g, zeta_new, lpw);
                double den = 0;                                                // delta . (lambda delta + b)
                for (int i = 0; i < N; i++) for (int k = 0; k < 6; k++) { const double d = delta[6 * (size_t)i + k]; den += d * (lambda * d + E.b[6 * (size_t)E.perm[i] + k]); }
                rho = (current - fresh) / (den + 1e-3);
                if (rho > 0) {
                    stop = stop || (current - fresh < c.min_relative_residual_increment * current);
                    if (stop) break;
                    double alpha = 1.0 - std::pow(2 * rho - 1, 3);
                    alpha = std::min(alpha, c.upper_scale_factor);
                    lambda *= std::max(c.lower_scale_factor, alpha);
                    ni = 2;
                    current = fresh;
                    zeta = zeta_new;
                    g.pose = trial.pose;
                    xnorm = pose_vector_norm();
                    update_confidence(g, zeta, lpw);
                    if (!compute_system(g, zeta, E)) return B2R_ERR_SINGULAR;
                    stop = stop || right_term_small();
                    if (stop) break;
                } else {
                    lambda *= ni; ni *= 2;
                }
            }
            lm_count++; rep.lm_steps++;
            stop = stop || lm_count >= c.max_iteration_lm;
        } while (!((rho > 0) || stop));
        rep.iterations++;
        stop = stop || current < c.min_residual || iter >= c.max_iteration;
    }
    rep.residual = current; rep.lambda = lambda;
    return B2R_OK;
}

// ValidatePoseGraphConnectivity: breadth-first over the edges that count
bool connected(const Graph& g, bool ignore_uncertain) {
    const int N = g.n();
    if (N == 0) return true;
    std::vector<std::vector<int>> adj(N);
    for (const Edge& e : g.edges) if (!ignore_uncertain || !e.uncertain) { adj[e.s].push_back(e.t); adj[e.t].push_back(e.s); }
    std::vector<char> seen(N, 0);
    std::queue<int> q; q.push(0); seen[0] = 1; int cnt = 1;
    while (!q.empty()) { const int u = q.front(); q.pop(); for (int v : adj[u]) if (!seen[v]) { seen[v] = 1; cnt++; q.push(v); } }
    return cnt == N;
}

}  // namespace

// o3d.pipelines.registration.global_optimization (Levenberg-Marquardt).  All matrices row-major f64, HOST memory.
//   poses[n_nodes][16] in/out; edges: src/dst ids, transformation[n_edges][16], information[n_edges][36],
//   uncertain[n_edges], confidence[n_edges] in/out; keep[n_edges] out: 1 = the edge is in the optimised graph
//   (upstream REPLACES the graph by the pruned one; the caller drops the edges with keep == 0).
//   option4 = {max_correspondence_distance, edge_prune_threshold, preference_loop_closure, reference_node};
//   criteria8 = {max_iteration, min_relative_increment, min_relative_residual_increment, min_right_term, min_residual,
//                max_iteration_lm, upper_scale_factor, lower_scale_factor};
//   report8 (may be NULL) = {valid, iterations pass 1, LM steps pass 1, iterations pass 2, LM steps pass 2,
//                            residual before, residual after pass 2, edges kept}.
// Returns B2R_OK also when upstream would only warn and leave the graph untouched (report8[0] = 0 then).
B2R_API int b2r_posegraph_optimize(int n_nodes, double* poses, int n_edges, const int32_t* src, const int32_t* dst,
                                   const double* transformation, const double* information, const uint8_t* uncertain,
                                   double* confidence, uint8_t* keep, const double* option4, const double* criteria8,
                                   double* report8) {
    if (n_nodes < 0 || n_edges < 0 || (n_nodes && !poses) || !option4 || !criteria8 ||
        (n_edges && (!src || !dst || !transformation || !information || !uncertain || !confidence || !keep))) {
        b2r_set_error("b2r_posegraph_optimize: bad argument");
        return B2R_ERR_BAD_ARG;
    }
    if (report8) for (int i = 0; i < 8; i++) report8[i] = 0.0;
    for (int k = 0; k < n_edges; k++) keep[k] = 1;
    Graph g;
    g.pose.assign(poses, poses + 16 * (size_t)n_nodes);
    g.edges.resize(n_edges);
    bool ids_ok = true;
    for (int k = 0; k < n_edges; k++) {
        Edge& e = g.edges[k];
        e.s = src[k]; e.t = dst[k];
        memcpy(e.T, transformation + 16 * (size_t)k, sizeof(e.T));
        memcpy(e.info, information + 36 * (size_t)k, sizeof(e.info));
        e.uncertain = uncertain[k] ? 1 : 0; e.conf = confidence[k];
        if (e.s < 0 || e.s >= n_nodes || e.t < 0 || e.t >= n_nodes) ids_ok = false;
    }
    // ValidatePoseGraph: connected, edge ids in range, certain edges carry confidence 1 -- otherwise upstream warns and returns
    if (!ids_ok) { b2r_set_error("Invalid PoseGraph - an edge references an invalid node."); return B2R_OK; }
    if (!connected(g, false)) { b2r_set_error("Invalid PoseGraph - graph is not connected."); return B2R_OK; }
    for (const Edge& e : g.edges) if (!e.uncertain && e.conf != 1.0) { b2r_set_error("Invalid PoseGraph - the certain edge does not have 1.0 as a confidence."); return B2R_OK; }
    const Option o = {option4[0], option4[1], option4[2], (int)option4[3]};
    const Criteria c = {(int)criteria8[0], criteria8[1], criteria8[2], criteria8[3], criteria8[4], (int)criteria8[5], criteria8[6], criteria8[7]};
    Report r1, r2;
    int rc = optimize_lm(g, c, o, r1);
    if (rc != B2R_OK) return rc;
    // CreatePoseGraphWithoutInvalidEdges
    Graph pruned;
    pruned.pose = g.pose;
    std::vector<int> origin;
    for (int k = 0; k < n_edges; k++) {
        const Edge& e = g.edges[k];
        if (!e.uncertain || e.conf > o.prune) { pruned.edges.push_back(e); origin.push_back(k); } else keep[k] = 0;
    }
    rc = optimize_lm(pruned, c, o, r2);
    if (rc != B2R_OK) return rc;
    // CompensateReferencePoseGraphNode
    if (o.reference >= 0 && o.reference < n_nodes) {
        double inv[16], comp[16];
        if (!b2r::m4_inverse(&pruned.pose[16 * (size_t)o.reference], inv)) { b2r_set_error("global_optimization: singular pose"); return B2R_ERR_SINGULAR; }
        b2r::m4_mul(poses + 16 * (size_t)o.reference, inv, comp);
        for (int i = 0; i < n_nodes; i++) b2r::m4_mul(comp, &pruned.pose[16 * (size_t)i], &pruned.pose[16 * (size_t)i]);
    }
    memcpy(poses, pruned.pose.data(), sizeof(double) * 16 * (size_t)n_nodes);
    for (size_t j = 0; j < origin.size(); j++) confidence[origin[j]] = pruned.edges[j].conf;
    for (int k = 0; k < n_edges; k++) if (!keep[k]) confidence[k] = g.edges[k].conf;
    if (report8) {
        report8[0] = 1; report8[1] = r1.iterations; report8[2] = r1.lm_steps; report8[3] = r2.iterations; report8[4] = r2.lm_steps;
        report8[5] = r1.residual0; report8[6] = r2.residual; report8[7] = (double)origin.size();
    }
    return B2R_OK;
}
