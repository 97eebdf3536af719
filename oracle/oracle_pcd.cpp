// oracle_pcd.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (same rules as oracle.cpp: only tests/, smoke() and the
// CPU legs of the benchmarks may load liboracle.so).
//
// CPU restatement of the point-cloud conditioning calls the reference scripts make on either side of ICP
// (SURVEY.md 8(f) row f1), as Open3D 0.19.0's legacy geometry::PointCloud implements them:
//   remove_statistical_outlier(20, 2.0)         CodeThatWorks/GoodColorReconstruction.py:144,
//                                               CodeThatWorks/BetterVisualizer2CamWorksCOLOR.py:531
//   voxel_down_sample(voxel)                    GoodColorReconstruction.py:147, BetterVisualizer2CamWorksCOLOR.py:643
//   estimate_normals(KDTreeSearchParamHybrid)   GoodColorReconstruction.py:150, BetterVisualizer2CamWorksCOLOR.py:537
//   orient_normals_towards_camera_location      GoodColorReconstruction.py:151
// PARITY UNPINNED (as for oracle.cpp): the upstream C++ is not in /root/reference; the semantics are restated from the
// published algorithms (PointCloud::VoxelDownSample / RemoveStatisticalOutliers, EstimateNormals.cpp: cumulant
// covariance + the closed-form symmetric 3x3 eigen solver of Eberly's "A Robust Eigensolver for 3x3 Symmetric
// Matrices").  Neighbour search is BRUTE FORCE here (obviously exact); ties in distance go to the lower index.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <tuple>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#define O3R_API extern "C" __attribute__((visibility("default")))

namespace {

struct Nb { double d2; int32_t i; };
inline bool nb_less(const Nb& a, const Nb& b) { return a.d2 < b.d2 || (a.d2 == b.d2 && a.i < b.i); }

// k nearest (radius <= 0) or hybrid: within radius (d2 <= r*r is NOT what nanoflann's radius search keeps: it keeps
// d2 < r*r ... upstream SearchHybrid passes radius*radius to a kNN result set bounded by that worst distance, which
// admits d2 < r2 only) and at most k, ascending (d2, index).
void search(const double* pts, int64_t n, const double* q, int k, double radius, std::vector<Nb>& out) {
    out.clear();
    const double r2 = radius > 0.0 ? radius * radius : INFINITY;
    std::vector<Nb> all;
    all.reserve(256);
    for (int64_t j = 0; j < n; j++) {
        const double dx = pts[3 * j] - q[0], dy = pts[3 * j + 1] - q[1], dz = pts[3 * j + 2] - q[2];
        const double d2 = dx * dx + dy * dy + dz * dz;
        if (d2 < r2) all.push_back({d2, (int32_t)j});
    }
    const size_t kk = std::min<size_t>((size_t)k, all.size());
    std::partial_sort(all.begin(), all.begin() + kk, all.end(), nb_less);
    out.assign(all.begin(), all.begin() + kk);
}

struct V3 { double x, y, z; };
inline V3 cross(const V3& a, const V3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline double dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// A = {a00, a01, a02, a11, a12, a22}
V3 eigenvector0(const double* A, double e) {
    const V3 r0{A[0] - e, A[1], A[2]}, r1{A[1], A[3] - e, A[4]}, r2{A[2], A[4], A[5] - e};
    const V3 c01 = cross(r0, r1), c02 = cross(r0, r2), c12 = cross(r1, r2);
    const double d0 = dot(c01, c01), d1 = dot(c02, c02), d2 = dot(c12, c12);
    double dmax = d0; int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) { imax = 2; }
    if (imax == 0) { const double s = std::sqrt(d0); return {c01.x / s, c01.y / s, c01.z / s}; }
    if (imax == 1) { const double s = std::sqrt(d1); return {c02.x / s, c02.y / s, c02.z / s}; }
    const double s = std::sqrt(d2); return {c12.x / s, c12.y / s, c12.z / s};
}

V3 eigenvector1(const double* A, const V3& e0, double eval1) {
    V3 U;
    if (std::fabs(e0.x) > std::fabs(e0.y)) { const double inv = 1.0 / std::sqrt(e0.x * e0.x + e0.z * e0.z); U = {-e0.z * inv, 0.0, e0.x * inv}; }
    else { const double inv = 1.0 / std::sqrt(e0.y * e0.y + e0.z * e0.z); U = {0.0, e0.z * inv, -e0.y * inv}; }
    const V3 V = cross(e0, U);
    const V3 AU{A[0] * U.x + A[1] * U.y + A[2] * U.z, A[1] * U.x + A[3] * U.y + A[4] * U.z, A[2] * U.x + A[4] * U.y + A[5] * U.z};
    const V3 AV{A[0] * V.x + A[1] * V.y + A[2] * V.z, A[1] * V.x + A[3] * V.y + A[4] * V.z, A[2] * V.x + A[4] * V.y + A[5] * V.z};
    double m00 = dot(U, AU) - eval1, m01 = dot(U, AV), m11 = dot(V, AV) - eval1;
    const double a00 = std::fabs(m00), a01 = std::fabs(m01), a11 = std::fabs(m11);
    if (a00 >= a11) {
        if (std::max(a00, a01) > 0.0) {
            if (a00 >= a01) { m01 /= m00; m00 = 1.0 / std::sqrt(1.0 + m01 * m01); m01 *= m00; }
            else { m00 /= m01; m01 = 1.0 / std::sqrt(1.0 + m00 * m00); m00 *= m01; }
            return {m01 * U.x - m00 * V.x, m01 * U.y - m00 * V.y, m01 * U.z - m00 * V.z};
        }
        return U;
    }
    if (std::max(a11, a01) > 0.0) {
        if (a11 >= a01) { m01 /= m11; m11 = 1.0 / std::sqrt(1.0 + m01 * m01); m01 *= m11; }
        else { m11 /= m01; m01 = 1.0 / std::sqrt(1.0 + m11 * m11); m11 *= m01; }
        return {m11 * U.x - m01 * V.x, m11 * U.y - m01 * V.y, m11 * U.z - m01 * V.z};
    }
    return U;
}

// eigenvector of the smallest eigenvalue of the symmetric matrix C = {c00, c01, c02, c11, c12, c22}
V3 fast_eigen3x3(const double* C) {
    double A[6];
    double maxc = C[0];
    for (int i = 1; i < 6; i++) maxc = std::max(maxc, C[i]);
    if (maxc == 0.0) return {0.0, 0.0, 0.0};
    for (int i = 0; i < 6; i++) A[i] = C[i] / maxc;
    const double norm = A[1] * A[1] + A[2] * A[2] + A[4] * A[4];
    if (norm > 0.0) {
        const double q = (A[0] + A[3] + A[5]) / 3.0;
        const double b00 = A[0] - q, b11 = A[3] - q, b22 = A[5] - q;
        const double p = std::sqrt((b00 * b00 + b11 * b11 + b22 * b22 + norm * 2.0) / 6.0);
        const double c00 = b11 * b22 - A[4] * A[4];
        const double c01 = A[1] * b22 - A[4] * A[2];
        const double c02 = A[1] * A[4] - b11 * A[2];
        const double det = (b00 * c00 - A[1] * c01 + A[2] * c02) / (p * p * p);
        const double half_det = std::min(std::max(det * 0.5, -1.0), 1.0);
        const double angle = std::acos(half_det) / 3.0;
        const double two_thirds_pi = 2.09439510239319549;
        const double beta2 = std::cos(angle) * 2.0;
        const double beta0 = std::cos(angle + two_thirds_pi) * 2.0;
        const double beta1 = -(beta0 + beta2);
        const double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;
        if (half_det >= 0.0) {
            const V3 v2 = eigenvector0(A, e2);
            if (e2 < e0 && e2 < e1) return v2;
            const V3 v1 = eigenvector1(A, v2, e1);
            if (e1 < e0 && e1 < e2) return v1;
            return cross(v1, v2);
        }
        const V3 v0 = eigenvector0(A, e0);
        if (e0 < e1 && e0 < e2) return v0;
        const V3 v1 = eigenvector1(A, v0, e1);
        if (e1 < e0 && e1 < e2) return v1;
        return cross(v0, v1);
    }
    if (C[0] < C[3] && C[0] < C[5]) return {1.0, 0.0, 0.0};
    if (C[3] < C[0] && C[3] < C[5]) return {0.0, 1.0, 0.0};
    return {0.0, 0.0, 1.0};
}

}  // namespace

// exact neighbour search: idx[nq][k] (-1 padded), d2[nq][k], counts[nq]
O3R_API void o3r_pcd_search(const double* pts, int64_t n, const double* q, int64_t nq, int k, double radius, int32_t* idx,
                            double* d2, int32_t* counts) {
#pragma omp parallel
    {
        std::vector<Nb> nb;
#pragma omp for schedule(dynamic, 64)
        for (int64_t i = 0; i < nq; i++) {
            search(pts, n, q + 3 * i, k, radius, nb);
            counts[i] = (int32_t)nb.size();
            for (int j = 0; j < k; j++) {
                idx[i * k + j] = j < (int)nb.size() ? nb[j].i : -1;
                if (d2) d2[i * k + j] = j < (int)nb.size() ? nb[j].d2 : -1.0;
            }
        }
    }
}

// PointCloud::EstimateNormals(search_param, fast_normal_computation = true).  radius <= 0: KNN(k); else Hybrid(radius, k).
// normals: in/out [n][3]; has_normals: existing normals orient the result.  cov_out (optional) [n][6].
O3R_API void o3r_pcd_estimate_normals(const double* pts, int64_t n, int k, double radius, double* normals, int has_normals,
                                      double* cov_out) {
#pragma omp parallel
    {
        std::vector<Nb> nb;
#pragma omp for schedule(dynamic, 64)
        for (int64_t i = 0; i < n; i++) {
            search(pts, n, pts + 3 * i, k, radius, nb);
            double C[6] = {1.0, 0.0, 0.0, 1.0, 0.0, 1.0};        // identity when fewer than 3 neighbours
            if (nb.size() >= 3) {
                double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
                for (const Nb& e : nb) {
                    const double x = pts[3 * e.i], y = pts[3 * e.i + 1], z = pts[3 * e.i + 2];
                    c[0] += x; c[1] += y; c[2] += z;
                    c[3] += x * x; c[4] += x * y; c[5] += x * z; c[6] += y * y; c[7] += y * z; c[8] += z * z;
                }
                const double m = (double)nb.size();
                for (int j = 0; j < 9; j++) c[j] /= m;
                C[0] = c[3] - c[0] * c[0]; C[3] = c[6] - c[1] * c[1]; C[5] = c[8] - c[2] * c[2];
                C[1] = c[4] - c[0] * c[1]; C[2] = c[5] - c[0] * c[2]; C[4] = c[7] - c[1] * c[2];
            }
            if (cov_out) std::memcpy(cov_out + 6 * i, C, sizeof(C));
            V3 nrm = fast_eigen3x3(C);
            const double nn = std::sqrt(dot(nrm, nrm));
            const V3 old{normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]};
            if (nn == 0.0) nrm = has_normals ? old : V3{0.0, 0.0, 1.0};
            if (has_normals && dot(nrm, old) < 0.0) { nrm.x = -nrm.x; nrm.y = -nrm.y; nrm.z = -nrm.z; }
            normals[3 * i] = nrm.x; normals[3 * i + 1] = nrm.y; normals[3 * i + 2] = nrm.z;
        }
    }
}

// PointCloud::OrientNormalsTowardsCameraLocation
O3R_API void o3r_pcd_orient_normals(const double* pts, int64_t n, const double* cam, double* normals) {
    for (int64_t i = 0; i < n; i++) {
        V3 ref{cam[0] - pts[3 * i], cam[1] - pts[3 * i + 1], cam[2] - pts[3 * i + 2]};
        double* nm = normals + 3 * i;
        const double nn = std::sqrt(nm[0] * nm[0] + nm[1] * nm[1] + nm[2] * nm[2]);
        if (nn == 0.0) {
            const double rn = std::sqrt(dot(ref, ref));
            if (rn == 0.0) { nm[0] = 0.0; nm[1] = 0.0; nm[2] = 1.0; }
            else { nm[0] = ref.x / rn; nm[1] = ref.y / rn; nm[2] = ref.z / rn; }
        } else if (nm[0] * ref.x + nm[1] * ref.y + nm[2] * ref.z < 0.0) {
            nm[0] = -nm[0]; nm[1] = -nm[1]; nm[2] = -nm[2];
        }
    }
}

// PointCloud::RemoveStatisticalOutliers(nb_neighbors, std_ratio): keep[i] = 1 for inliers; avg (optional) = mean
// neighbour distance per point (the point itself is its own first neighbour, as upstream's KNN search returns it).
O3R_API int64_t o3r_pcd_statistical_outlier(const double* pts, int64_t n, int nb_neighbors, double std_ratio, uint8_t* keep,
                                            double* avg_out) {
    if (n == 0) return 0;
    std::vector<double> avg((size_t)n);
    int64_t valid = 0;
#pragma omp parallel
    {
        std::vector<Nb> nb;
#pragma omp for schedule(dynamic, 64) reduction(+ : valid)
        for (int64_t i = 0; i < n; i++) {
            search(pts, n, pts + 3 * i, nb_neighbors, -1.0, nb);
            double mean = -1.0;
            if (!nb.empty()) {
                valid++;
                double s = 0.0;
                for (const Nb& e : nb) s += std::sqrt(e.d2);
                mean = s / (double)nb.size();
            }
            avg[i] = mean;
        }
    }
    double sum = 0.0;
    for (int64_t i = 0; i < n; i++) if (avg[i] > 0.0) sum += avg[i];
    const double cloud_mean = sum / (double)valid;
    double sq = 0.0;
    for (int64_t i = 0; i < n; i++) if (avg[i] > 0.0) sq += (avg[i] - cloud_mean) * (avg[i] - cloud_mean);
    const double std_dev = std::sqrt(sq / (double)(valid - 1));
    const double thr = cloud_mean + std_ratio * std_dev;
    int64_t kept = 0;
    for (int64_t i = 0; i < n; i++) {
        keep[i] = (avg[i] > 0.0 && avg[i] < thr) ? 1 : 0;
        kept += keep[i];
        if (avg_out) avg_out[i] = avg[i];
    }
    return kept;
}

// PointCloud::VoxelDownSample(voxel).  Output in ascending voxel-index order (upstream's order is that of an
// unordered_map: implementation-defined; comparisons are on the canonical order).  col / nrm may be null.
// Returns the number of output points; out_* sized by the caller to n.  out_key (optional) [m][3].
O3R_API int64_t o3r_pcd_voxel_down_sample(const double* pts, const double* col, const double* nrm, int64_t n, double voxel,
                                          double* out_pts, double* out_col, double* out_nrm, int32_t* out_key) {
    if (n == 0) return 0;
    double mn[3] = {pts[0], pts[1], pts[2]};
    for (int64_t i = 1; i < n; i++)
        for (int a = 0; a < 3; a++) mn[a] = std::min(mn[a], pts[3 * i + a]);
    for (int a = 0; a < 3; a++) mn[a] -= voxel * 0.5;
    struct Acc { double p[3] = {0, 0, 0}, c[3] = {0, 0, 0}, nn[3] = {0, 0, 0}; int64_t num = 0; };
    std::map<std::tuple<int, int, int>, Acc> vox;
    for (int64_t i = 0; i < n; i++) {
        int k[3];
        for (int a = 0; a < 3; a++) k[a] = (int)std::floor((pts[3 * i + a] - mn[a]) / voxel);
        Acc& A = vox[std::make_tuple(k[0], k[1], k[2])];
        for (int a = 0; a < 3; a++) A.p[a] += pts[3 * i + a];
        if (nrm && !std::isnan(nrm[3 * i]) && !std::isnan(nrm[3 * i + 1]) && !std::isnan(nrm[3 * i + 2]))
            for (int a = 0; a < 3; a++) A.nn[a] += nrm[3 * i + a];
        if (col) for (int a = 0; a < 3; a++) A.c[a] += col[3 * i + a];
        A.num++;
    }
    int64_t m = 0;
    for (const auto& kv : vox) {
        const Acc& A = kv.second;
        const double d = (double)A.num;
        for (int a = 0; a < 3; a++) {
            out_pts[3 * m + a] = A.p[a] / d;
            if (col) out_col[3 * m + a] = A.c[a] / d;
            if (nrm) out_nrm[3 * m + a] = A.nn[a] / d;
        }
        if (out_key) { out_key[3 * m] = std::get<0>(kv.first); out_key[3 * m + 1] = std::get<1>(kv.first); out_key[3 * m + 2] = std::get<2>(kv.first); }
        m++;
    }
    return m;
}

// ---- f4: compute_fpfh_feature (BetterVisualizer2CamWorksCOLOR.py:211-214) ------------------------------------------------
// Open3D 0.19 pipelines/registration/Feature.cpp: ComputePairFeatures / ComputeSPFHFeature / ComputeFPFHFeature, with PCL's
// weighting (sum of neighbour SPFHs divided by squared distance, normalised to 100 per angle, plus the point's own SPFH).
// One deviation, stated here and mirrored by the CUDA kernel: upstream picks the frame origin with
// acos(|a1|) > acos(|a2|); acos is strictly decreasing, so this restates it as |a1| < |a2| (no libm call on the decision).
namespace {
void pair_features(const double* p1, const double* n1, const double* p2, const double* n2, double f[3]) {
    f[0] = f[1] = f[2] = 0.0;
    double dp[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
    const double d = std::sqrt(dp[0] * dp[0] + dp[1] * dp[1] + dp[2] * dp[2]);
    if (d == 0.0) return;
    const double a1 = (n1[0] * dp[0] + n1[1] * dp[1] + n1[2] * dp[2]) / d;
    const double a2 = (n2[0] * dp[0] + n2[1] * dp[1] + n2[2] * dp[2]) / d;
    const double* na = n1;
    const double* nb = n2;
    double phi = a1;
    if (std::fabs(a1) < std::fabs(a2)) {
        na = n2; nb = n1;
        for (int c = 0; c < 3; c++) dp[c] = -dp[c];
        phi = -a2;
    }
    V3 v = cross(V3{dp[0], dp[1], dp[2]}, V3{na[0], na[1], na[2]});
    const double vn = std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z);
    if (vn == 0.0) return;
    v = {v.x / vn, v.y / vn, v.z / vn};
    const V3 w = cross(V3{na[0], na[1], na[2]}, v);
    f[2] = phi;
    f[1] = v.x * nb[0] + v.y * nb[1] + v.z * nb[2];
    f[0] = std::atan2(w.x * nb[0] + w.y * nb[1] + w.z * nb[2], na[0] * nb[0] + na[1] * nb[1] + na[2] * nb[2]);
}
inline int clamp_bin(double x) {
    int h = (int)std::floor(x);
    return h < 0 ? 0 : (h >= 11 ? 10 : h);
}
}  // namespace

// out [n][33] point-major.  radius <= 0: KNN(max_nn); else Hybrid(radius, max_nn).  Returns -1 without normals.
O3R_API int o3r_compute_fpfh(const double* pts, const double* nrm, int64_t n, int max_nn, double radius, double* out) {
    if (!nrm) return -1;
    const double PI = 3.14159265358979323846;
    std::vector<double> spfh((size_t)n * 33, 0.0);
    std::vector<std::vector<Nb>> nbs((size_t)n);
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; i++) {
        search(pts, n, pts + 3 * i, max_nn, radius, nbs[i]);
        const std::vector<Nb>& nb = nbs[i];
        if (nb.size() > 1) {
            const double inc = 100.0 / (double)(nb.size() - 1);
            for (size_t k = 1; k < nb.size(); k++) {
                double f[3];
                pair_features(pts + 3 * i, nrm + 3 * i, pts + 3 * (int64_t)nb[k].i, nrm + 3 * (int64_t)nb[k].i, f);
                spfh[i * 33 + clamp_bin(11 * (f[0] + PI) / (2.0 * PI))] += inc;
                spfh[i * 33 + 11 + clamp_bin(11 * (f[1] + 1.0) * 0.5)] += inc;
                spfh[i * 33 + 22 + clamp_bin(11 * (f[2] + 1.0) * 0.5)] += inc;
            }
        }
    }
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; i++) {
        double* o = out + i * 33;
        for (int j = 0; j < 33; j++) o[j] = 0.0;
        const std::vector<Nb>& nb = nbs[i];
        if (nb.size() <= 1) continue;
        double sum[3] = {0.0, 0.0, 0.0};
        for (size_t k = 1; k < nb.size(); k++) {
            const double dist = nb[k].d2;
            if (dist == 0.0) continue;
            for (int j = 0; j < 33; j++) {
                const double val = spfh[(size_t)nb[k].i * 33 + j] / dist;
                sum[j / 11] += val;
                o[j] += val;
            }
        }
        for (int j = 0; j < 3; j++) if (sum[j] != 0.0) sum[j] = 100.0 / sum[j];
        for (int j = 0; j < 33; j++) { o[j] *= sum[j / 11]; o[j] += spfh[i * 33 + j]; }
    }
    return 0;
}

// ---- f4: registration_ransac_based_on_feature_matching (BetterVisualizer2CamWorksCOLOR.py:230-243) ------------------------
// Open3D 0.19 Registration.cpp: nearest-feature correspondences, then RegistrationRANSACBasedOnCorrespondence.  Upstream's draw
// order depends on the OpenMP schedule and a per-thread mt19937, so the project fixes ONE deterministic algorithm with the same
// distribution (stated in csrc/features.cu and restated here, sequentially): splitmix-style counter generator per (iteration,
// draw), Horn's closed-form rigid fit (upstream: Eigen::umeyama -- same optimum), upstream's checkers, full validation by
// nearest neighbour, upstream's IsBetterRANSACThan and confidence rule, survivors taken in iteration order.
namespace {
inline uint64_t splitmix(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}
inline uint32_t draw(uint64_t seed, uint64_t itr, int j, uint32_t C) {
    const uint64_t h = splitmix(seed + itr * 0x9E3779B97F4A7C15ull + (uint64_t)(j + 1) * 0xD1B54A32D192ED03ull);
    return (uint32_t)(((h >> 32) * (uint64_t)C) >> 32);
}
// Horn 1987 (closed-form absolute orientation with unit quaternions); cyclic Jacobi on the symmetric 4x4
void horn(const double (*s)[3], const double (*t)[3], int n, double* T) {
    double ms[3] = {0, 0, 0}, mt[3] = {0, 0, 0};
    for (int i = 0; i < n; i++) for (int c = 0; c < 3; c++) { ms[c] += s[i][c]; mt[c] += t[i][c]; }
    for (int c = 0; c < 3; c++) { ms[c] /= (double)n; mt[c] /= (double)n; }
    double S[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < n; i++) for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) S[a][b] += (s[i][a] - ms[a]) * (t[i][b] - mt[b]);
    double A[4][4], V[4][4];
    A[0][0] = S[0][0] + S[1][1] + S[2][2];
    A[1][1] = S[0][0] - S[1][1] - S[2][2];
    A[2][2] = S[1][1] - S[0][0] - S[2][2];
    A[3][3] = S[2][2] - S[0][0] - S[1][1];
    A[0][1] = A[1][0] = S[1][2] - S[2][1];
    A[0][2] = A[2][0] = S[2][0] - S[0][2];
    A[0][3] = A[3][0] = S[0][1] - S[1][0];
    A[1][2] = A[2][1] = S[0][1] + S[1][0];
    A[1][3] = A[3][1] = S[2][0] + S[0][2];
    A[2][3] = A[3][2] = S[1][2] + S[2][1];
    for (int a = 0; a < 4; a++) for (int b = 0; b < 4; b++) V[a][b] = a == b ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 16; sweep++) {
        double off = 0.0;
        for (int p = 0; p < 3; p++) for (int q = p + 1; q < 4; q++) off += A[p][q] * A[p][q];
        if (off == 0.0) break;
        for (int p = 0; p < 3; p++)
            for (int q = p + 1; q < 4; q++) {
                const double apq = A[p][q];
                if (apq == 0.0) continue;
                const double theta = (A[q][q] - A[p][p]) / (2.0 * apq);
                const double tt = (theta >= 0.0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(tt * tt + 1.0), sn = tt * c;
                for (int k = 0; k < 4; k++) { const double x = A[k][p], y = A[k][q]; A[k][p] = c * x - sn * y; A[k][q] = sn * x + c * y; }
                for (int k = 0; k < 4; k++) { const double x = A[p][k], y = A[q][k]; A[p][k] = c * x - sn * y; A[q][k] = sn * x + c * y; }
                for (int k = 0; k < 4; k++) { const double x = V[k][p], y = V[k][q]; V[k][p] = c * x - sn * y; V[k][q] = sn * x + c * y; }
            }
    }
    int best = 0;
    for (int k = 1; k < 4; k++) if (A[k][k] > A[best][best]) best = k;
    double w = V[0][best], x = V[1][best], y = V[2][best], z = V[3][best];
    const double qn = std::sqrt(w * w + x * x + y * y + z * z);
    w /= qn; x /= qn; y /= qn; z /= qn;
    const double R[3][3] = {{1.0 - 2.0 * (y * y + z * z), 2.0 * (x * y - w * z), 2.0 * (x * z + w * y)},
                            {2.0 * (x * y + w * z), 1.0 - 2.0 * (x * x + z * z), 2.0 * (y * z - w * x)},
                            {2.0 * (x * z - w * y), 2.0 * (y * z + w * x), 1.0 - 2.0 * (x * x + y * y)}};
    for (int a = 0; a < 3; a++) {
        for (int b = 0; b < 3; b++) T[4 * a + b] = R[a][b];
        T[4 * a + 3] = mt[a] - (R[a][0] * ms[0] + R[a][1] * ms[1] + R[a][2] * ms[2]);
    }
    T[12] = T[13] = T[14] = 0.0; T[15] = 1.0;
}
inline void apply(const double* T, const double* p, double* o) {
    for (int r = 0; r < 3; r++) o[r] = T[4 * r] * p[0] + T[4 * r + 1] * p[1] + T[4 * r + 2] * p[2] + T[4 * r + 3];
}
}  // namespace

O3R_API void o3r_horn_rigid(const double* s, const double* t, int n, double* T) {
    horn(reinterpret_cast<const double (*)[3]>(s), reinterpret_cast<const double (*)[3]>(t), n, T);
}

// nearest feature by squared L2 (index order summation, ties -> lower index)
O3R_API void o3r_match_features(const double* qf, int64_t nq, const double* tf, int64_t nt, int dim, int32_t* out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nq; i++) {
        double best = INFINITY; int32_t bi = -1;
        for (int64_t j = 0; j < nt; j++) {
            double d2 = 0.0;
            for (int d = 0; d < dim; d++) { const double e = qf[i * dim + d] - tf[j * dim + d]; d2 += e * e; }
            if (d2 < best) { best = d2; bi = (int32_t)j; }
        }
        out[i] = bi;
    }
}

// report4 = {iterations consumed, hypotheses validated, winning iteration, its inlier-correspondence count}
O3R_API int o3r_ransac_correspondence(const double* src, int64_t N, const double* tgt, int64_t M, const int32_t* corres, int64_t C,
                                      double max_dist, int ransac_n, double edge_thr, double dist_thr, int64_t max_iteration,
                                      double confidence, uint64_t seed, double* out_T, double* out_fitness, double* out_rmse, int64_t* report4) {
    for (int i = 0; i < 16; i++) out_T[i] = (i % 5 == 0) ? 1.0 : 0.0;
    *out_fitness = 0.0; *out_rmse = 0.0;
    for (int i = 0; i < 4; i++) report4[i] = 0;
    if (ransac_n < 3 || C < ransac_n || !(max_dist > 0.0)) return 0;
    if (ransac_n > 8) return -1;
    confidence = std::min(std::max(confidence, 0.0), 1.0);
    const double r2 = max_dist * max_dist;
    double best_fit = 0.0, best_rmse = 0.0;
    int64_t est_k = max_iteration, validated = 0, consumed = 0, win = -1, win_inl = 0;
    for (int64_t itr = 0; itr < max_iteration && itr < est_k; itr++) {
        consumed = itr + 1;
        double s[8][3], t[8][3];
        for (int j = 0; j < ransac_n; j++) {
            const uint32_t c = draw(seed, (uint64_t)itr, j, (uint32_t)C);
            for (int k = 0; k < 3; k++) { s[j][k] = src[3 * (int64_t)corres[2 * c] + k]; t[j][k] = tgt[3 * (int64_t)corres[2 * c + 1] + k]; }
        }
        bool ok = true;
        if (edge_thr > 0.0)
            for (int i = 0; i < ransac_n && ok; i++)
                for (int j = i + 1; j < ransac_n && ok; j++) {
                    const double ds = std::sqrt((s[i][0] - s[j][0]) * (s[i][0] - s[j][0]) + (s[i][1] - s[j][1]) * (s[i][1] - s[j][1]) + (s[i][2] - s[j][2]) * (s[i][2] - s[j][2]));
                    const double dt = std::sqrt((t[i][0] - t[j][0]) * (t[i][0] - t[j][0]) + (t[i][1] - t[j][1]) * (t[i][1] - t[j][1]) + (t[i][2] - t[j][2]) * (t[i][2] - t[j][2]));
                    if (ds < dt * edge_thr || dt < ds * edge_thr) ok = false;
                }
        if (!ok) continue;
        double T[16];
        horn(s, t, ransac_n, T);
        if (dist_thr > 0.0)
            for (int i = 0; i < ransac_n && ok; i++) {
                double p[3]; apply(T, s[i], p);
                const double d2 = (t[i][0] - p[0]) * (t[i][0] - p[0]) + (t[i][1] - p[1]) * (t[i][1] - p[1]) + (t[i][2] - p[2]) * (t[i][2] - p[2]);
                if (std::sqrt(d2) > dist_thr) ok = false;
            }
        for (int k = 0; k < 16; k++) if (!std::isfinite(T[k])) ok = false;
        if (!ok) continue;
        validated++;
        int64_t cnt = 0; double sum = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : cnt, sum)
        for (int64_t i = 0; i < N; i++) {
            double p[3]; apply(T, src + 3 * i, p);
            double bd = INFINITY;
            for (int64_t j = 0; j < M; j++) {
                const double dx = tgt[3 * j] - p[0], dy = tgt[3 * j + 1] - p[1], dz = tgt[3 * j + 2] - p[2];
                const double d2 = dx * dx + dy * dy + dz * dz;
                if (d2 < bd) bd = d2;
            }
            if (bd < r2) { cnt++; sum += bd; }
        }
        const double fit = (double)cnt / (double)N, rm = cnt > 0 ? std::sqrt(sum / (double)cnt) : 0.0;
        if (fit > best_fit || (fit == best_fit && rm < best_rmse)) {
            best_fit = fit; best_rmse = rm; win = itr; std::memcpy(out_T, T, sizeof(T));
            int64_t inl = 0;
            for (int64_t c = 0; c < C; c++) {
                double p[3]; apply(T, src + 3 * (int64_t)corres[2 * c], p);
                const double* q = tgt + 3 * (int64_t)corres[2 * c + 1];
                const double d2 = (q[0] - p[0]) * (q[0] - p[0]) + (q[1] - p[1]) * (q[1] - p[1]) + (q[2] - p[2]) * (q[2] - p[2]);
                if (std::sqrt(d2) < max_dist) inl++;
            }
            win_inl = inl;
            const double ratio = (double)inl / (double)C;
            const double k_d = std::log(1.0 - confidence) / std::log(1.0 - std::pow(ratio, (double)ransac_n));
            if (std::isfinite(k_d) && k_d >= 0.0 && k_d < (double)est_k) est_k = (int64_t)std::ceil(k_d);
        }
    }
    if (consumed > est_k) consumed = est_k;
    *out_fitness = best_fit; *out_rmse = best_rmse;
    report4[0] = consumed; report4[1] = validated; report4[2] = win; report4[3] = win_inl;
    return 0;
}
