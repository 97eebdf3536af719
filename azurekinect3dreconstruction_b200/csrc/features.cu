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
#include <math.h>

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
