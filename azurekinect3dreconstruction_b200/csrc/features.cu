// features.cu -- SURVEY 8(f) row f4: FPFH features and feature-matching RANSAC for auto-extrinsics.
//
// Replaces, behind the C-ABI in include/b2recon.h,
//   o3d.pipelines.registration.compute_fpfh_feature(pcd_down, KDTreeSearchParamHybrid(radius=voxel*5, max_nn=100))
//   o3d.pipelines.registration.registration_ransac_based_on_feature_matching(src, tgt, src_fpfh, tgt_fpfh, ...)
// (/root/reference/CodeThatWorks/BetterVisualizer2CamWorksCOLOR.py:197-243).  Semantics: Open3D 0.19
// pipelines/registration/Feature.cpp (Rusu et al. 2009, PCL's weighting) and Registration.cpp.
//
// FPFH: one exact hybrid neighbour search through the uniform grid of pointcloud.cu (b2r_knn_search, neighbours ascending
// by (distance, index) as KDTreeFlann returns them); SPFH: one thread per point walks its neighbours and bins the three
// Darboux-frame angles in f64; FPFH: one WARP per point, lane j owns histogram bin j, so a neighbour's 33-bin SPFH row is
// one coalesced 264-byte read; the per-group normalisers are summed in upstream's order (neighbour-major, bin-minor)
// by lanes 0 / 11 / 22 through shuffles, which keeps the result bit-identical to the sequential CPU loop.
#include "common.cuh"
#include "host_math.h"
#include "rigid_fit.cuh"
#include <math.h>
#include <vector>
#include <algorithm>
#include <string.h>

struct b2r_knn;
extern "C" b2r_knn* b2r_knn_create(const double* xyz, int64_t n, double cell, int k_hint, cudaStream_t stream);
extern "C" void b2r_knn_destroy(b2r_knn* h);
extern "C" int b2r_knn_search(const b2r_knn* h, const double* query, int64_t nq, int k, double radius, int32_t* idx, double* d2,
                              int32_t* counts, cudaStream_t stream);

namespace {

#define FPFH_PI 3.14159265358979323846

struct F3 { double x, y, z; };
__device__ __forceinline__ double f3dot(const F3& a, const F3& b) { return xdadd(xdadd(xdmul(a.x, b.x), xdmul(a.y, b.y)), xdmul(a.z, b.z)); }
__device__ __forceinline__ F3 f3cross(const F3& a, const F3& b) {
    return {xdsub(xdmul(a.y, b.z), xdmul(a.z, b.y)), xdsub(xdmul(a.z, b.x), xdmul(a.x, b.z)), xdsub(xdmul(a.x, b.y), xdmul(a.y, b.x))};
}

// ComputePairFeatures (Feature.cpp): f[0] = theta, f[1] = alpha, f[2] = phi.  The frame is anchored at the point whose
// normal makes the smaller angle with the connecting line: acos(|a1|) > acos(|a2|)  <=>  |a1| < |a2|.
__device__ void pair_features(const F3& p1, const F3& n1, const F3& p2, const F3& n2, double* f) {
    f[0] = f[1] = f[2] = 0.0;
    F3 dp{xdsub(p2.x, p1.x), xdsub(p2.y, p1.y), xdsub(p2.z, p1.z)};
    const double d = sqrt(f3dot(dp, dp));
    if (d == 0.0) return;
    const double a1 = xddiv(f3dot(n1, dp), d), a2 = xddiv(f3dot(n2, dp), d);
    F3 na = n1, nb = n2;
    double phi;
    if (fabs(a1) < fabs(a2)) { na = n2; nb = n1; dp = {-dp.x, -dp.y, -dp.z}; phi = -a2; } else phi = a1;
    F3 v = f3cross(dp, na);
    const double vn = sqrt(f3dot(v, v));
    if (vn == 0.0) return;
    v = {xddiv(v.x, vn), xddiv(v.y, vn), xddiv(v.z, vn)};
    const F3 w = f3cross(na, v);
    f[2] = phi;
    f[1] = f3dot(v, nb);
    f[0] = atan2(f3dot(w, nb), f3dot(na, nb));
}

__device__ __forceinline__ int bin11(double x) {
    int h = (int)floor(x);
    return h < 0 ? 0 : (h >= 11 ? 10 : h);
}

// ComputeSPFHFeature: spfh[i][33], point-major
__global__ void __launch_bounds__(128) spfh_kernel(const double* __restrict__ xyz, const double* __restrict__ nrm, const int32_t* __restrict__ nbr,
                                                   const int32_t* __restrict__ counts, int k, int64_t n, double* __restrict__ spfh) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double hist[33];
#pragma unroll
    for (int j = 0; j < 33; j++) hist[j] = 0.0;
    const int cnt = counts[i];
    if (cnt > 1) {
        const F3 p{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]}, nn{nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]};
        const double inc = xddiv(100.0, (double)(cnt - 1));
        for (int q = 1; q < cnt; q++) {
            const int64_t j = nbr[i * k + q];
            const F3 p2{xyz[3 * j], xyz[3 * j + 1], xyz[3 * j + 2]}, n2{nrm[3 * j], nrm[3 * j + 1], nrm[3 * j + 2]};
            double f[3];
            pair_features(p, nn, p2, n2, f);
            const int h0 = bin11(xddiv(xdmul(11.0, xdadd(f[0], FPFH_PI)), xdmul(2.0, FPFH_PI)));
            const int h1 = bin11(xdmul(xdmul(11.0, xdadd(f[1], 1.0)), 0.5));
            const int h2 = bin11(xdmul(xdmul(11.0, xdadd(f[2], 1.0)), 0.5));
            hist[h0] = xdadd(hist[h0], inc);
            hist[11 + h1] = xdadd(hist[11 + h1], inc);
            hist[22 + h2] = xdadd(hist[22 + h2], inc);
        }
    }
    for (int j = 0; j < 33; j++) spfh[i * 33 + j] = hist[j];
}

// ComputeFPFHFeature, second pass: one warp per point.  Lane l owns bin l; lane 0 also owns bin 32 (slot B).
__global__ void __launch_bounds__(128) fpfh_kernel(const double* __restrict__ spfh, const int32_t* __restrict__ nbr, const double* __restrict__ d2,
                                                   const int32_t* __restrict__ counts, int k, int64_t n, double* __restrict__ out) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= n) return;
    const int cnt = counts[i];
    double accA = 0.0, accB = 0.0;           // bins lane and 32
    double gsum = 0.0;                       // lanes 0, 11, 22: normaliser of group lane / 11
    if (cnt > 1) {
        for (int q = 1; q < cnt; q++) {
            const double dist = d2[i * k + q];
            if (dist == 0.0) continue;       // warp-uniform
            const int64_t j = nbr[i * k + q];
            const double vA = xddiv(spfh[j * 33 + lane], dist);
            const double vB = xddiv(spfh[j * 33 + 32], dist);          // every lane computes it; only group 2's sum and lane 0 use it
            accA = xdadd(accA, vA);
            accB = xdadd(accB, vB);
            // sum[g] += val in bin order within the group (bins 11 g .. 11 g + 10); bin 32 lives in vB
#pragma unroll
            for (int t = 0; t < 11; t++) {
                const int src = (lane / 11) * 11 + t;                  // lanes 0..10 -> group 0, 11..21 -> 1, 22..31 -> 2
                const double v = __shfl_sync(0xffffffffu, vA, src < 32 ? src : 31);
                const double add = (src < 32) ? v : vB;                // group 2, t = 10: bin 32
                if (lane == 0 || lane == 11 || lane == 22) gsum = xdadd(gsum, add);
            }
        }
    }
    // broadcast the normalisers
    double s0 = __shfl_sync(0xffffffffu, gsum, 0), s1 = __shfl_sync(0xffffffffu, gsum, 11), s2 = __shfl_sync(0xffffffffu, gsum, 22);
    if (s0 != 0.0) s0 = xddiv(100.0, s0);
    if (s1 != 0.0) s1 = xddiv(100.0, s1);
    if (s2 != 0.0) s2 = xddiv(100.0, s2);
    if (cnt > 1) {
        const double sc = lane < 11 ? s0 : (lane < 22 ? s1 : s2);
        out[i * 33 + lane] = xdadd(xdmul(accA, sc), spfh[i * 33 + lane]);
        if (lane == 0) out[i * 33 + 32] = xdadd(xdmul(accB, s2), spfh[i * 33 + 32]);
    } else {
        out[i * 33 + lane] = 0.0;
        if (lane == 0) out[i * 33 + 32] = 0.0;
    }
}

}  // namespace

// compute_fpfh_feature(pcd, KDTreeSearchParamHybrid(radius, max_nn)) (radius <= 0: KDTreeSearchParamKNN(max_nn)).
// DEVICE f64: xyz, normals [n][3]; out [n][33] point-major (upstream's Feature.data is its transpose, 33 x n).
B2R_API int b2r_compute_fpfh(const double* xyz, const double* normals, int64_t n, int max_nn, double radius, double* out,
                             cudaStream_t stream) {
    if (!normals) { b2r_set_error("[ComputeFPFHFeature] Failed because input point cloud has no normal."); return B2R_ERR_FORMAT; }
    if (!xyz || !out || n < 0 || max_nn < 1 || max_nn > 128 || n > 0x7fffffff) { b2r_set_error("b2r_compute_fpfh: bad argument (1 <= max_nn <= 128)"); return B2R_ERR_BAD_ARG; }
    if (n == 0) return B2R_OK;
    int32_t *nbr = nullptr, *cnt = nullptr;
    double *d2 = nullptr, *spfh = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&nbr, (size_t)n * max_nn * 4, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&cnt, (size_t)n * 4, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&d2, (size_t)n * max_nn * 8, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&spfh, (size_t)n * 33 * 8, stream);
    int rc = B2R_OK;
    if (e != cudaSuccess) { b2r_set_error("b2r_compute_fpfh: %s", cudaGetErrorString(e)); rc = B2R_ERR_CUDA; }
    b2r_knn* tree = nullptr;
    if (rc == B2R_OK) {
        tree = b2r_knn_create(xyz, n, radius > 0.0 ? radius : 0.0, max_nn, stream);
        if (!tree) rc = B2R_ERR_CUDA;
    }
    if (rc == B2R_OK) rc = b2r_knn_search(tree, nullptr, n, max_nn, radius, nbr, d2, cnt, stream);
    if (rc == B2R_OK) {
        spfh_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(xyz, normals, nbr, cnt, max_nn, n, spfh);
        fpfh_kernel<<<(unsigned)((n * 32 + 127) / 128), 128, 0, stream>>>(spfh, nbr, d2, cnt, max_nn, n, out);
        if (cudaGetLastError() != cudaSuccess) { b2r_set_error("b2r_compute_fpfh: launch failed"); rc = B2R_ERR_CUDA; }
    }
    if (tree) b2r_knn_destroy(tree);
    cudaFreeAsync(nbr, stream); cudaFreeAsync(cnt, stream); cudaFreeAsync(d2, stream); cudaFreeAsync(spfh, stream);
    return rc;
}

// ================================================================================================================
// registration_ransac_based_on_feature_matching / _on_correspondence (Registration.cpp)
// ================================================================================================================
// Upstream draws ransac_n correspondences per iteration from a per-thread std::mt19937, estimates a rigid transform
// (Eigen::umeyama), runs the cheap checkers, and fully validates the survivors (nearest target point of every transformed
// source point).  Its result depends on the OpenMP schedule; the algorithm is fixed here as a deterministic one with the
// same distribution: iteration `itr` draws index j as mulhi(mix64(seed + itr * K1 + j * K2) >> 32, C) -- a counter-based
// generator, so a batch of 65 536 iterations is one kernel launch; survivors are validated by one CTA each
// (b2r_knn_evaluate_transforms) and then scanned IN ITERATION ORDER on the host with upstream's rules (strictly better
// fitness, then smaller rmse; confidence-based shrinking of the iteration budget).  The rotation comes from Horn's closed
// form (largest eigenvector of the 4x4 quaternion matrix, cyclic Jacobi) instead of an SVD: same optimum, proper rotation by
// construction, and only + - * / sqrt, so the CPU restatement reproduces it bit for bit.
extern "C" int b2r_knn_evaluate_transforms(const b2r_knn* h, const double* src_xyz, int64_t N, const double* T, int64_t H, double max_dist,
                                           int64_t* count, double* sumd2, cudaStream_t stream);

namespace {

#define RANSAC_MAX_N 8
#define RANSAC_BATCH 65536
#define RANSAC_K1 0x9E3779B97F4A7C15ull
#define RANSAC_K2 0xD1B54A32D192ED03ull

__host__ __device__ inline uint32_t ransac_draw(uint64_t seed, uint64_t itr, int j, uint32_t C) {
    const uint64_t h = b2r_mix64(seed + itr * RANSAC_K1 + (uint64_t)(j + 1) * RANSAC_K2);
    return (uint32_t)(((h >> 32) * (uint64_t)C) >> 32);
}

struct RansacArgs {
    const double* src; const double* tgt;
    const int32_t* corres; uint32_t C;
    int n;                        // ransac_n
    double edge_thr, dist_thr;    // <= 0: checker absent
    uint64_t seed, itr0; int count;
    int32_t* surv_itr; double* surv_T; int* surv_n;
};

__global__ void __launch_bounds__(128) ransac_hypothesis_kernel(const RansacArgs a) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= a.count) return;
    const uint64_t itr = a.itr0 + (uint64_t)tid;
    double s[RANSAC_MAX_N][3], t[RANSAC_MAX_N][3];
    for (int j = 0; j < a.n; j++) {
        const uint32_t c = ransac_draw(a.seed, itr, j, a.C);
        const int si = a.corres[2 * c], ti = a.corres[2 * c + 1];
        for (int k = 0; k < 3; k++) { s[j][k] = a.src[3 * (int64_t)si + k]; t[j][k] = a.tgt[3 * (int64_t)ti + k]; }
    }
    if (a.edge_thr > 0.0) {       // CorrespondenceCheckerBasedOnEdgeLength
        for (int i = 0; i < a.n; i++)
            for (int j = i + 1; j < a.n; j++) {
                const double dsx = xdsub(s[i][0], s[j][0]), dsy = xdsub(s[i][1], s[j][1]), dsz = xdsub(s[i][2], s[j][2]);
                const double dtx = xdsub(t[i][0], t[j][0]), dty = xdsub(t[i][1], t[j][1]), dtz = xdsub(t[i][2], t[j][2]);
                const double ds = sqrt(xdadd(xdadd(xdmul(dsx, dsx), xdmul(dsy, dsy)), xdmul(dsz, dsz)));
                const double dt = sqrt(xdadd(xdadd(xdmul(dtx, dtx), xdmul(dty, dty)), xdmul(dtz, dtz)));
                if (ds < xdmul(dt, a.edge_thr) || dt < xdmul(ds, a.edge_thr)) return;
            }
    }
    double T[16];
    horn_rigid(s, t, a.n, T);
    if (a.dist_thr > 0.0) {       // CorrespondenceCheckerBasedOnDistance
        for (int i = 0; i < a.n; i++) {
            double d2 = 0.0;
            for (int r = 0; r < 3; r++) {
                const double p = xdadd(xdadd(xdadd(xdmul(T[4 * r], s[i][0]), xdmul(T[4 * r + 1], s[i][1])), xdmul(T[4 * r + 2], s[i][2])), T[4 * r + 3]);
                const double e = xdsub(t[i][r], p);
                d2 = xdadd(d2, xdmul(e, e));
            }
            if (sqrt(d2) > a.dist_thr) return;
        }
    }
    for (int k = 0; k < 16; k++) if (!isfinite(T[k])) return;
    const int slot = atomicAdd(a.surv_n, 1);
    a.surv_itr[slot] = tid;
    for (int k = 0; k < 16; k++) a.surv_T[(int64_t)slot * 16 + k] = T[k];
}

// EvaluateInlierCorrespondenceRatio of every survivor: correspondences (i, j) with |T s_i - t_j| < max_dist
__global__ void __launch_bounds__(256) ransac_corres_inlier_kernel(const double* __restrict__ src, const double* __restrict__ tgt, const int32_t* __restrict__ corres,
                                                                   uint32_t C, const double* __restrict__ T, double max_dist, int32_t* __restrict__ out) {
    __shared__ double sT[16];
    __shared__ int s_c[8];
    if (threadIdx.x < 16) sT[threadIdx.x] = T[(int64_t)blockIdx.x * 16 + threadIdx.x];
    __syncthreads();
    int cnt = 0;
    for (uint32_t c = threadIdx.x; c < C; c += blockDim.x) {
        const double* s = src + 3 * (int64_t)corres[2 * c];
        const double* t = tgt + 3 * (int64_t)corres[2 * c + 1];
        double d2 = 0.0;
        for (int r = 0; r < 3; r++) {
            const double p = xdadd(xdadd(xdadd(xdmul(sT[4 * r], s[0]), xdmul(sT[4 * r + 1], s[1])), xdmul(sT[4 * r + 2], s[2])), sT[4 * r + 3]);
            const double e = xdsub(t[r], p);
            d2 = xdadd(d2, xdmul(e, e));
        }
        if (sqrt(d2) < max_dist) cnt++;
    }
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, o);
    if ((threadIdx.x & 31) == 0) s_c[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) { int c = 0; for (int w = 0; w < 8; w++) c += s_c[w]; out[blockIdx.x] = c; }
}

__global__ void ransac_transform_kernel(const double* __restrict__ src, int64_t N, const double* __restrict__ T, double* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const double x = src[3 * i], y = src[3 * i + 1], z = src[3 * i + 2];
    double h[4];
#pragma unroll
    for (int r = 0; r < 4; r++) h[r] = xdadd(xdadd(xdadd(xdmul(T[4 * r], x), xdmul(T[4 * r + 1], y)), xdmul(T[4 * r + 2], z)), T[4 * r + 3]);
    out[3 * i] = xddiv(h[0], h[3]); out[3 * i + 1] = xddiv(h[1], h[3]); out[3 * i + 2] = xddiv(h[2], h[3]);
}

// nearest feature (squared L2 over `dim` values summed in index order; ties -> lower index): corres_ij of upstream
__global__ void __launch_bounds__(128) feature_match_kernel(const double* __restrict__ qf, int64_t nq, const double* __restrict__ tf, int64_t nt, int dim,
                                                            int32_t* __restrict__ out) {
    extern __shared__ double tile[];                       // 64 target features
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double q[40];
    if (i < nq) for (int d = 0; d < dim; d++) q[d] = qf[i * dim + d];
    double best = INFINITY; int32_t bi = -1;
    for (int64_t t0 = 0; t0 < nt; t0 += 64) {
        const int m = (int)min((int64_t)64, nt - t0);
        __syncthreads();
        for (int e = threadIdx.x; e < m * dim; e += blockDim.x) tile[e] = tf[t0 * dim + e];
        __syncthreads();
        if (i < nq)
            for (int j = 0; j < m; j++) {
                double d2 = 0.0;
                for (int d = 0; d < dim; d++) { const double e = xdsub(q[d], tile[j * dim + d]); d2 = xdadd(d2, xdmul(e, e)); }
                if (d2 < best) { best = d2; bi = (int32_t)(t0 + j); }
            }
    }
    if (i < nq) out[i] = bi;
}

}  // namespace

// corres_ij of RegistrationRANSACBasedOnFeatureMatching: for every query feature the index of the nearest target feature.
// DEVICE f64 point-major [n][dim], dim <= 40.
B2R_API int b2r_match_features(const double* query_feat, int64_t nq, const double* target_feat, int64_t nt, int dim, int32_t* out_idx,
                               cudaStream_t stream) {
    if (!query_feat || !target_feat || !out_idx || nq < 0 || nt < 1 || dim < 1 || dim > 40) { b2r_set_error("b2r_match_features: bad argument (1 <= dim <= 40)"); return B2R_ERR_BAD_ARG; }
    if (nq == 0) return B2R_OK;
    feature_match_kernel<<<(unsigned)((nq + 127) / 128), 128, 64 * dim * sizeof(double), stream>>>(query_feat, nq, target_feat, nt, dim, out_idx);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// registration_ransac_based_on_correspondence(source, target, corres, max_dist, TransformationEstimationPointToPoint(false),
// ransac_n, [EdgeLength(edge_thr), Distance(dist_thr)], RANSACConvergenceCriteria(max_iteration, confidence)).
// DEVICE: src [N][3], tgt [M][3], corres int32 [C][2].  HOST outputs; report4 (optional) = {iterations consumed, hypotheses
// validated, index of the winning iteration, its inlier-correspondence count}.  No acceptable hypothesis: identity, fitness 0.
B2R_API int b2r_ransac_correspondence(const double* src_xyz, int64_t N, const double* tgt_xyz, int64_t M, const int32_t* corres, int64_t C,
                                      double max_dist, int ransac_n, double edge_thr, double dist_thr, int64_t max_iteration,
                                      double confidence, uint64_t seed, double* out_T_host, double* fitness_host, double* rmse_host,
                                      int64_t* report4_host, int32_t* corr_out, cudaStream_t stream) {
    if (!out_T_host || !fitness_host || !rmse_host) return B2R_ERR_BAD_ARG;
    b2r::m4_identity(out_T_host);
    *fitness_host = 0.0; *rmse_host = 0.0;
    if (report4_host) for (int i = 0; i < 4; i++) report4_host[i] = 0;
    // upstream returns an empty RegistrationResult for these
    if (ransac_n < 3 || C < ransac_n || !(max_dist > 0.0)) return B2R_OK;
    if (ransac_n > RANSAC_MAX_N || !src_xyz || !tgt_xyz || !corres || N < 1 || M < 1 || C > 0x7fffffff || max_iteration < 0) {
        b2r_set_error("b2r_ransac_correspondence: bad argument (3 <= ransac_n <= 8)");
        return B2R_ERR_BAD_ARG;
    }
    confidence = confidence < 0.0 ? 0.0 : (confidence > 1.0 ? 1.0 : confidence);
    b2r_knn* tree = b2r_knn_create(tgt_xyz, M, max_dist, 1, stream);
    if (!tree) return B2R_ERR_CUDA;
    int32_t *surv_itr = nullptr, *inl = nullptr; double *surv_T = nullptr, *sumd2 = nullptr; int64_t* cnt = nullptr; int* surv_n = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&surv_itr, RANSAC_BATCH * 4, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&inl, RANSAC_BATCH * 4, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&surv_T, (size_t)RANSAC_BATCH * 128, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&sumd2, RANSAC_BATCH * 8, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&cnt, RANSAC_BATCH * 8, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&surv_n, 4, stream);
    int rc = B2R_OK;
    if (e != cudaSuccess) { b2r_set_error("b2r_ransac_correspondence: %s", cudaGetErrorString(e)); rc = B2R_ERR_CUDA; }
    std::vector<int32_t> h_itr(RANSAC_BATCH), h_inl(RANSAC_BATCH), order;
    std::vector<int64_t> h_cnt(RANSAC_BATCH);
    std::vector<double> h_sum(RANSAC_BATCH), h_T((size_t)RANSAC_BATCH * 16);
    double best_fit = 0.0, best_rmse = 0.0;
    bool have = false;
    int64_t est_k = max_iteration, validated = 0, consumed = 0, win_itr = -1, win_inl = 0;
    for (int64_t itr0 = 0; rc == B2R_OK && itr0 < max_iteration && itr0 < est_k; itr0 += RANSAC_BATCH) {
        const int count = (int)std::min<int64_t>(RANSAC_BATCH, std::min(max_iteration, est_k) - itr0);
        RansacArgs a = {src_xyz, tgt_xyz, corres, (uint32_t)C, ransac_n, edge_thr, dist_thr, seed, (uint64_t)itr0, count, surv_itr, surv_T, surv_n};
        if (cudaMemsetAsync(surv_n, 0, 4, stream) != cudaSuccess) { rc = B2R_ERR_CUDA; break; }
        ransac_hypothesis_kernel<<<(count + 127) / 128, 128, 0, stream>>>(a);
        int ns = 0;
        if (cudaMemcpyAsync(&ns, surv_n, 4, cudaMemcpyDeviceToHost, stream) != cudaSuccess || cudaStreamSynchronize(stream) != cudaSuccess) { b2r_set_error("b2r_ransac_correspondence: hypothesis kernel failed"); rc = B2R_ERR_CUDA; break; }
        consumed = itr0 + count;
        if (ns == 0) continue;
        rc = b2r_knn_evaluate_transforms(tree, src_xyz, N, surv_T, ns, max_dist, cnt, sumd2, stream);
        if (rc != B2R_OK) break;
        ransac_corres_inlier_kernel<<<ns, 256, 0, stream>>>(src_xyz, tgt_xyz, corres, (uint32_t)C, surv_T, max_dist, inl);
        cudaMemcpyAsync(h_itr.data(), surv_itr, (size_t)ns * 4, cudaMemcpyDeviceToHost, stream);
        cudaMemcpyAsync(h_inl.data(), inl, (size_t)ns * 4, cudaMemcpyDeviceToHost, stream);
        cudaMemcpyAsync(h_cnt.data(), cnt, (size_t)ns * 8, cudaMemcpyDeviceToHost, stream);
        cudaMemcpyAsync(h_sum.data(), sumd2, (size_t)ns * 8, cudaMemcpyDeviceToHost, stream);
        cudaMemcpyAsync(h_T.data(), surv_T, (size_t)ns * 128, cudaMemcpyDeviceToHost, stream);
        if (cudaStreamSynchronize(stream) != cudaSuccess) { b2r_set_error("b2r_ransac_correspondence: validation failed"); rc = B2R_ERR_CUDA; break; }
        order.resize(ns);
        for (int i = 0; i < ns; i++) order[i] = i;
        std::sort(order.begin(), order.end(), [&](int x, int y) { return h_itr[x] < h_itr[y]; });
        for (int oi = 0; oi < ns; oi++) {
            const int sidx = order[oi];
            const int64_t itr = itr0 + h_itr[sidx];
            if (itr >= est_k) { consumed = est_k; break; }
            validated++;
            const double fit = (double)h_cnt[sidx] / (double)N;
            const double rm = h_cnt[sidx] > 0 ? std::sqrt(h_sum[sidx] / (double)h_cnt[sidx]) : 0.0;
            // IsBetterRANSACThan (the initial best is the empty result: fitness 0, rmse 0)
            if (fit > best_fit || (fit == best_fit && rm < best_rmse)) {
                best_fit = fit; best_rmse = rm; have = true; win_itr = itr; win_inl = h_inl[sidx];
                memcpy(out_T_host, &h_T[(size_t)sidx * 16], 128);
                const double ratio = (double)h_inl[sidx] / (double)C;
                const double k_d = std::log(1.0 - confidence) / std::log(1.0 - std::pow(ratio, (double)ransac_n));
                if (std::isfinite(k_d) && k_d >= 0.0 && k_d < (double)est_k) est_k = (int64_t)std::ceil(k_d);   // ratio 0 or confidence 1: no shrinking
            }
        }
    }
    if (consumed > est_k) consumed = est_k;
    if (rc == B2R_OK && corr_out) {       // correspondence_set of the winner: nearest target of every transformed source point (-1: none)
        double* moved = nullptr; int32_t* c1 = nullptr; double* Td = nullptr;
        cudaError_t e2 = cudaMallocAsync((void**)&moved, (size_t)N * 24, stream);
        if (e2 == cudaSuccess) e2 = cudaMallocAsync((void**)&c1, (size_t)N * 4, stream);
        if (e2 == cudaSuccess) e2 = cudaMallocAsync((void**)&Td, 128, stream);
        if (e2 == cudaSuccess) e2 = cudaMemcpyAsync(Td, out_T_host, 128, cudaMemcpyHostToDevice, stream);
        if (e2 == cudaSuccess) {
            ransac_transform_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(src_xyz, N, Td, moved);
            rc = b2r_knn_search(tree, moved, N, 1, max_dist, corr_out, nullptr, c1, stream);
            if (rc == B2R_OK && cudaStreamSynchronize(stream) != cudaSuccess) rc = B2R_ERR_CUDA;
        } else { b2r_set_error("b2r_ransac_correspondence: %s", cudaGetErrorString(e2)); rc = B2R_ERR_CUDA; }
        cudaFreeAsync(moved, stream); cudaFreeAsync(c1, stream); cudaFreeAsync(Td, stream);
    }
    b2r_knn_destroy(tree);
    cudaFreeAsync(surv_itr, stream); cudaFreeAsync(inl, stream); cudaFreeAsync(surv_T, stream); cudaFreeAsync(sumd2, stream); cudaFreeAsync(cnt, stream); cudaFreeAsync(surv_n, stream);
    if (rc != B2R_OK) return rc;
    if (have) { *fitness_host = best_fit; *rmse_host = best_rmse; }
    if (report4_host) { report4_host[0] = consumed; report4_host[1] = validated; report4_host[2] = win_itr; report4_host[3] = win_inl; }
    return B2R_OK;
}
