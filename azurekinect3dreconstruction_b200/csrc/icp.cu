// icp.cu -- point-to-plane ICP on sm_100a (north_star subsystem 4, second half).
//
// Replaces o3d.pipelines.registration.registration_icp(src, tgt, max_dist, init,
//   TransformationEstimationPointToPlane(), ICPConvergenceCriteria(...))
// (/root/reference/CodeThatWorks/GoodColorReconstruction.py:194-198,
//  CodeThatWorks/BetterVisualizer2CamWorksCOLOR.py:255-259, azure-kinect-DK/Reconstruction/Reconstruction.py:129-133);
// upstream semantics restated in SURVEY.md Appendix A.7.
//
// Design: upstream builds a FLANN kd-tree and loops on the host.  Here the target is binned once into a
// uniform grid (cell = max_dist; cub radix sort of cell ids), and ONE kernel per ICP iteration does
//   apply the previous update U to every source point -> exact 1-NN within max_dist over the <=27
//   neighbouring cells -> residual + Jacobian row -> f64 reduction of JtJ/Jtr/err^2/count
//   (warp shuffles -> shared memory -> per-block partials) -> the last block solves the 6x6 system,
//   accumulates T <- U*T, evaluates fitness / inlier_rmse and the convergence test, all on the device.
// The host only reads the final state.
#include "common.cuh"
#include "host_math.h"
#include "rigid_fit.cuh"
#include <cub/cub.cuh>
#include <math.h>
#include <mutex>

namespace {

#define ICP_THREADS 256
#define ICP_NV 30     // 21 JtJ + 6 Jtr + r2 + err2 + count

struct IcpGrid {
    double lo[3];
    double cell;
    int dim[3];
};

struct IcpState {
    double T[16];          // accumulated transformation
    double U[16];          // update to apply to the working cloud at the next launch
    double fitness, rmse, prev_fitness, prev_rmse;
    long long n_corr;
    int done, iters, launches;
    unsigned ticket;
    double bounds[6];      // target min xyz, max xyz
};

__global__ void __launch_bounds__(1024) bounds_kernel(const double* __restrict__ p, long long M, IcpState* st) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (long long i = threadIdx.x; i < M; i += blockDim.x)
#pragma unroll
        for (int k = 0; k < 3; k++) { const double v = p[3 * i + k]; lo[k] = fmin(lo[k], v); hi[k] = fmax(hi[k], v); }
    __shared__ double s[32][6];
#pragma unroll
    for (int k = 0; k < 3; k++)
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fmin(lo[k], __shfl_down_sync(0xffffffffu, lo[k], o));
            hi[k] = fmax(hi[k], __shfl_down_sync(0xffffffffu, hi[k], o));
        }
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 3; k++) { s[threadIdx.x >> 5][k] = lo[k]; s[threadIdx.x >> 5][3 + k] = hi[k]; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); w++)
            for (int k = 0; k < 3; k++) { s[0][k] = fmin(s[0][k], s[w][k]); s[0][3 + k] = fmax(s[0][3 + k], s[w][3 + k]); }
        for (int k = 0; k < 6; k++) st->bounds[k] = s[0][k];
    }
}

__device__ __forceinline__ int grid_coord(const IcpGrid& g, double v, int k) {
    const int c = (int)floor(xddiv(xdsub(v, g.lo[k]), g.cell));
    return min(max(c, 0), g.dim[k] - 1);
}

__global__ void cell_id_kernel(const double* __restrict__ p, int M, const IcpGrid g, unsigned* cid, int* idx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    const int a = grid_coord(g, p[3 * i], 0), b = grid_coord(g, p[3 * i + 1], 1), c = grid_coord(g, p[3 * i + 2], 2);
    cid[i] = (unsigned)((a * g.dim[1] + b) * g.dim[2] + c);
    idx[i] = i;
}

__global__ void cell_range_kernel(const unsigned* __restrict__ cid_sorted, int M, int* cell_start, int* cell_end) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    const unsigned c = cid_sorted[i];
    if (i == 0 || cid_sorted[i - 1] != c) cell_start[c] = i;
    if (i == M - 1 || cid_sorted[i + 1] != c) cell_end[c] = i + 1;
}

__global__ void copy_points_kernel(const double* __restrict__ in, double* out, long long n3) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n3) out[i] = in[i];
}

// Every loop is fully unrolled and the pivot row / column exchange is written as predicated swaps over STATIC indices, so
// the 36 + 36 + 24 doubles live in registers instead of dynamically indexed local memory (this runs on ONE thread at the
// end of every Gauss-Newton iteration: its latency is on the critical path every ICP iteration).  Same operations, same
// order, same results as the loop form the oracle uses.
__device__ bool icp_ldlt6(const double* A_in, const double* b_in, double* x) {
    constexpr int n = 6;
    double A[36], L[36], D[6], y[6], z[6];
    int perm[6];
#pragma unroll
    for (int i = 0; i < 36; i++) { A[i] = A_in[i]; L[i] = 0.0; }
#pragma unroll
    for (int i = 0; i < n; i++) { perm[i] = i; D[i] = 0.0; }
    bool stopped = false;
#pragma unroll
    for (int k = 0; k < n; k++) {
        if (stopped) continue;
        int p = k;
        double best = fabs(A[k * n + k]);
#pragma unroll
        for (int i = k + 1; i < n; i++) if (fabs(A[i * n + i]) > best) { best = fabs(A[i * n + i]); p = i; }
#pragma unroll
        for (int i = k + 1; i < n; i++) {
            if (i == p) {
#pragma unroll
                for (int j = 0; j < n; j++) { const double t = A[k * n + j]; A[k * n + j] = A[i * n + j]; A[i * n + j] = t; }
#pragma unroll
                for (int j = 0; j < n; j++) { const double t = A[j * n + k]; A[j * n + k] = A[j * n + i]; A[j * n + i] = t; }
#pragma unroll
                for (int j = 0; j < k; j++) { const double t = L[k * n + j]; L[k * n + j] = L[i * n + j]; L[i * n + j] = t; }
                const int t = perm[k]; perm[k] = perm[i]; perm[i] = t;
            }
        }
        const double d = A[k * n + k];
        D[k] = d;
        L[k * n + k] = 1.0;
        if (d == 0.0) {
#pragma unroll
            for (int i = k + 1; i < n; i++) { D[i] = 0.0; L[i * n + i] = 1.0; L[i * n + k] = 0.0; }
            stopped = true;
            continue;
        }
#pragma unroll
        for (int i = k + 1; i < n; i++) L[i * n + k] = xddiv(A[i * n + k], d);
#pragma unroll
        for (int i = k + 1; i < n; i++)
#pragma unroll
            for (int j = k + 1; j < n; j++) A[i * n + j] = xdsub(A[i * n + j], xdmul(xdmul(L[i * n + k], d), L[j * n + k]));
    }
#pragma unroll
    for (int i = 0; i < n; i++) {
        double v = 0.0;
#pragma unroll
        for (int j = 0; j < n; j++) if (perm[i] == j) v = b_in[j];
        y[i] = v;
    }
#pragma unroll
    for (int i = 0; i < n; i++) {
        double s = y[i];
#pragma unroll
        for (int j = 0; j < i; j++) s = xdsub(s, xdmul(L[i * n + j], z[j]));
        z[i] = s;
    }
#pragma unroll
    for (int i = 0; i < n; i++) z[i] = (fabs(D[i]) > 2.2250738585072014e-308) ? xddiv(z[i], D[i]) : 0.0;
#pragma unroll
    for (int i = n - 1; i >= 0; i--) {
        double s = z[i];
#pragma unroll
        for (int j = i + 1; j < n; j++) s = xdsub(s, xdmul(L[j * n + i], y[j]));
        y[i] = s;
    }
    bool ok = true;
#pragma unroll
    for (int i = 0; i < n; i++)
#pragma unroll
        for (int j = 0; j < n; j++) if (perm[i] == j) x[j] = y[i];
#pragma unroll
    for (int i = 0; i < n; i++) ok = ok && isfinite(x[i]);
    return ok;
}

__device__ void icp_vec6_to_m4(const double* x, double* T) {
    const double ca = cos(x[0]), sa = sin(x[0]), cb = cos(x[1]), sb = sin(x[1]), cg = cos(x[2]), sg = sin(x[2]);
    T[0] = cg * cb; T[1] = cg * sb * sa - sg * ca; T[2] = cg * sb * ca + sg * sa; T[3] = x[3];
    T[4] = sg * cb; T[5] = sg * sb * sa + cg * ca; T[6] = sg * sb * ca - cg * sa; T[7] = x[4];
    T[8] = -sb;     T[9] = cb * sa;                T[10] = cb * ca;               T[11] = x[5];
    T[12] = 0; T[13] = 0; T[14] = 0; T[15] = 1;
}

}  // namespace

namespace {

struct IcpArgs {
    double* pcd;                 // N*3 working copy of the source (transformed in place)
    long long N;
    const double* tgt;           // M*3
    const double* nrm;           // M*3
    const int* sorted_idx;       // M, target indices ordered by cell
    const int* cell_start;
    const int* cell_end;
    IcpGrid g;
    double max_dist;
    int32_t* corr;               // N
    double* partials;
    IcpState* st;
    int max_iter;
    double rel_fitness, rel_rmse;
    // TransformationEstimationForColoredICP (f4): two residual rows per correspondence
    int colored;
    const double* src_int;       // N: (r + g + b) / 3 of the source colours
    const double* tgt_int;       // M
    const double* tgt_grad;      // M*3: colour gradient on the target's tangent planes
    double sqrt_lg, sqrt_lp;     // sqrt(lambda_geometric), sqrt(1 - lambda_geometric)
    int p2p;                     // TransformationEstimationPointToPoint: accumulate the cross-covariance sums instead of JtJ / Jtr
};

__global__ void __launch_bounds__(ICP_THREADS) icp_iter_kernel(const IcpArgs a) {
    IcpState* st = a.st;
    if (st->done) return;
    __shared__ double sU[16];
    if (threadIdx.x < 16) sU[threadIdx.x] = st->U[threadIdx.x];
    __syncthreads();
    double acc[ICP_NV];
#pragma unroll
    for (int i = 0; i < ICP_NV; i++) acc[i] = 0.0;
    const double r2max = xdmul(a.max_dist, a.max_dist);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.N; i += (long long)gridDim.x * blockDim.x) {
        // pcd.Transform(U)  (Geometry3D::Transform: homogeneous divide)
        const double x = a.pcd[3 * i], y = a.pcd[3 * i + 1], z = a.pcd[3 * i + 2];
        double h[4];
#pragma unroll
        for (int r = 0; r < 4; r++) h[r] = xdadd(xdadd(xdadd(xdmul(sU[4 * r], x), xdmul(sU[4 * r + 1], y)), xdmul(sU[4 * r + 2], z)), sU[4 * r + 3]);
        const double s[3] = {xddiv(h[0], h[3]), xddiv(h[1], h[3]), xddiv(h[2], h[3])};
        a.pcd[3 * i] = s[0]; a.pcd[3 * i + 1] = s[1]; a.pcd[3 * i + 2] = s[2];
        // exact nearest neighbour within max_dist
        int c0[3], c1[3];
        bool any = true;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            c0[k] = (int)floor(xddiv(xdsub(xdsub(s[k], a.max_dist), a.g.lo[k]), a.g.cell));
            c1[k] = (int)floor(xddiv(xdsub(xdadd(s[k], a.max_dist), a.g.lo[k]), a.g.cell));
            if (c1[k] < 0 || c0[k] >= a.g.dim[k]) any = false;
            c0[k] = max(c0[k], 0); c1[k] = min(c1[k], a.g.dim[k] - 1);
        }
        int best = -1;
        double bd = 0.0;
        if (any) {
            for (int ca = c0[0]; ca <= c1[0]; ca++)
                for (int cb = c0[1]; cb <= c1[1]; cb++)
                    for (int cc = c0[2]; cc <= c1[2]; cc++) {
                        const int ci = (ca * a.g.dim[1] + cb) * a.g.dim[2] + cc;
                        const int e = a.cell_end[ci];
                        for (int q = a.cell_start[ci]; q < e; q++) {
                            const int j = a.sorted_idx[q];
                            const double dx = xdsub(a.tgt[3 * j], s[0]), dy = xdsub(a.tgt[3 * j + 1], s[1]), dz = xdsub(a.tgt[3 * j + 2], s[2]);
                            const double d2 = xdadd(xdadd(xdmul(dx, dx), xdmul(dy, dy)), xdmul(dz, dz));
                            if (d2 > r2max) continue;
                            if (best < 0 || d2 < bd || (d2 == bd && j < best)) { bd = d2; best = j; }
                        }
                    }
        }
        a.corr[i] = best;
        if (best >= 0 && a.p2p) {
            // sums for the closed-form rigid fit, taken about the grid origin (keeps sum s t^T - n ms mt^T well conditioned)
            const double* t = a.tgt + 3 * best;
            const double sc[3] = {xdsub(s[0], a.g.lo[0]), xdsub(s[1], a.g.lo[1]), xdsub(s[2], a.g.lo[2])};
            const double tc[3] = {xdsub(t[0], a.g.lo[0]), xdsub(t[1], a.g.lo[1]), xdsub(t[2], a.g.lo[2])};
#pragma unroll
            for (int p = 0; p < 3; p++) {
#pragma unroll
                for (int q = 0; q < 3; q++) acc[3 * p + q] = xdadd(acc[3 * p + q], xdmul(sc[p], tc[q]));
                acc[9 + p] = xdadd(acc[9 + p], sc[p]);
                acc[12 + p] = xdadd(acc[12 + p], tc[p]);
            }
            acc[27] = xdadd(acc[27], bd);
            acc[28] = xdadd(acc[28], bd);
            acc[29] += 1.0;
        } else if (best >= 0) {
            const double* t = a.tgt + 3 * best;
            const double* n = a.nrm + 3 * best;
            double r = xdadd(xdadd(xdmul(xdsub(s[0], t[0]), n[0]), xdmul(xdsub(s[1], t[1]), n[1])), xdmul(xdsub(s[2], t[2]), n[2]));
            double J[6] = {xdsub(xdmul(s[1], n[2]), xdmul(s[2], n[1])), xdsub(xdmul(s[2], n[0]), xdmul(s[0], n[2])),
                           xdsub(xdmul(s[0], n[1]), xdmul(s[1], n[0])), n[0], n[1], n[2]};
            if (a.colored) {
                // photometric row: the source point projected into the target's tangent plane, intensity predicted by the
                // target's colour gradient there; d(is0_proj)/d(vs) = dit^T (I - n n^T)
                const double dn = r;                                        // (vs - vt) . nt
                const double* dit = a.tgt_grad + 3 * best;
                double proj[3], ditM[3];
#pragma unroll
                for (int q = 0; q < 3; q++) proj[q] = xdsub(xdsub(s[q], xdmul(dn, n[q])), t[q]);   // vs_proj - vt
                const double is0 = xdadd(xdadd(xdadd(xdmul(dit[0], proj[0]), xdmul(dit[1], proj[1])), xdmul(dit[2], proj[2])), a.tgt_int[best]);
                const double dd = xdadd(xdadd(xdmul(dit[0], n[0]), xdmul(dit[1], n[1])), xdmul(dit[2], n[2]));
#pragma unroll
                for (int q = 0; q < 3; q++) ditM[q] = xdmul(a.sqrt_lp, xdsub(xdmul(dd, n[q]), dit[q]));   // sqrt(lp) * (-dit^T M)
                const double r1 = xdmul(a.sqrt_lp, xdsub(a.src_int[i], is0));
                const double J1[6] = {xdsub(xdmul(s[1], ditM[2]), xdmul(s[2], ditM[1])), xdsub(xdmul(s[2], ditM[0]), xdmul(s[0], ditM[2])),
                                      xdsub(xdmul(s[0], ditM[1]), xdmul(s[1], ditM[0])), ditM[0], ditM[1], ditM[2]};
                int k1 = 0;
#pragma unroll
                for (int p = 0; p < 6; p++)
#pragma unroll
                    for (int q = p; q < 6; q++) { acc[k1] = xdadd(acc[k1], xdmul(J1[p], J1[q])); k1++; }
#pragma unroll
                for (int p = 0; p < 6; p++) acc[21 + p] = xdadd(acc[21 + p], xdmul(J1[p], r1));
                acc[27] = xdadd(acc[27], xdmul(r1, r1));
                r = xdmul(a.sqrt_lg, r);
#pragma unroll
                for (int p = 0; p < 6; p++) J[p] = xdmul(a.sqrt_lg, J[p]);
            }
            int k = 0;
#pragma unroll
            for (int p = 0; p < 6; p++)
#pragma unroll
                for (int q = p; q < 6; q++) { acc[k] = xdadd(acc[k], xdmul(J[p], J[q])); k++; }
#pragma unroll
            for (int p = 0; p < 6; p++) acc[21 + p] = xdadd(acc[21 + p], xdmul(J[p], r));
            acc[27] = xdadd(acc[27], xdmul(r, r));
            acc[28] = xdadd(acc[28], bd);
            acc[29] += 1.0;
        }
    }
    __shared__ double sh[ICP_THREADS / 32][ICP_NV];
    __shared__ bool is_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < ICP_NV; i++) {
        double v = acc[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = xdadd(v, __shfl_down_sync(0xffffffffu, v, o));
        if (lane == 0) sh[warp][i] = v;
    }
    __syncthreads();
    if (threadIdx.x < ICP_NV) {
        double v = 0.0;
        for (int w = 0; w < ICP_THREADS / 32; w++) v = xdadd(v, sh[w][threadIdx.x]);
        a.partials[(size_t)blockIdx.x * ICP_NV + threadIdx.x] = v;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(&st->ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (!is_last) return;
    // fixed tree over the blocks' partials: warp w sums blocks w, w + 8, ... (four loads in flight), then the warp sums in order
    __shared__ double tot[ICP_NV];
    __shared__ double part[ICP_THREADS / 32][ICP_NV];
    if (lane < ICP_NV) {
        constexpr int NW = ICP_THREADS / 32;
        double v0 = 0.0, v1 = 0.0, v2 = 0.0, v3 = 0.0;
        unsigned b = warp;
        for (; b + 3 * NW < gridDim.x; b += 4 * NW) {
            const double a0 = __ldcg(&a.partials[(size_t)b * ICP_NV + lane]), a1 = __ldcg(&a.partials[(size_t)(b + NW) * ICP_NV + lane]);
            const double a2 = __ldcg(&a.partials[(size_t)(b + 2 * NW) * ICP_NV + lane]), a3 = __ldcg(&a.partials[(size_t)(b + 3 * NW) * ICP_NV + lane]);
            v0 = xdadd(v0, a0); v1 = xdadd(v1, a1); v2 = xdadd(v2, a2); v3 = xdadd(v3, a3);
        }
        for (; b < gridDim.x; b += NW) v0 = xdadd(v0, __ldcg(&a.partials[(size_t)b * ICP_NV + lane]));
        part[warp][lane] = xdadd(xdadd(v0, v1), xdadd(v2, v3));
    }
    __syncthreads();
    if (threadIdx.x < ICP_NV) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < ICP_THREADS / 32; w++) v = xdadd(v, part[w][threadIdx.x]);
        tot[threadIdx.x] = v;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    st->ticket = 0;
    const int k = st->launches;          // 0 = initial Evaluate, k >= 1 = k-th loop iteration
    st->launches = k + 1;
    const double n = tot[29];
    const double fitness = a.N ? xddiv(n, (double)a.N) : 0.0;
    const double rmse = n > 0.0 ? sqrt(xddiv(tot[28], n)) : 0.0;
    if (k >= 1) {
        double R[16];
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) { double s = 0.0; for (int q = 0; q < 4; q++) s = xdadd(s, xdmul(sU[i * 4 + q], st->T[q * 4 + j])); R[i * 4 + j] = s; }
        for (int i = 0; i < 16; i++) st->T[i] = R[i];
        st->iters = k;
    }
    st->prev_fitness = st->fitness; st->prev_rmse = st->rmse;
    st->fitness = fitness; st->rmse = rmse; st->n_corr = (long long)n;
    bool done = false;
    if (k >= 1 && fabs(xdsub(st->prev_fitness, fitness)) < a.rel_fitness && fabs(xdsub(st->prev_rmse, rmse)) < a.rel_rmse) done = true;
    if (k >= a.max_iter) done = true;
    if (done) { st->done = 1; return; }
    // TransformationEstimationPointToPlane::ComputeTransformation for the next iteration
    double U[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    if (n > 0.0 && a.p2p) {
        // TransformationEstimationPointToPoint::ComputeTransformation (no scaling): closed-form rigid fit of the pairs
        double msr[3], mtr[3], ms[3], mt[3], S[3][3];
        for (int p = 0; p < 3; p++) { msr[p] = xddiv(tot[9 + p], n); mtr[p] = xddiv(tot[12 + p], n); }
        for (int p = 0; p < 3; p++)
            for (int q = 0; q < 3; q++) S[p][q] = xdsub(tot[3 * p + q], xdmul(xdmul(n, msr[p]), mtr[q]));
        for (int p = 0; p < 3; p++) { ms[p] = xdadd(msr[p], a.g.lo[p]); mt[p] = xdadd(mtr[p], a.g.lo[p]); }
        double Tn[16];
        horn_from_covariance(S, ms, mt, Tn);
        bool fin = true;
        for (int i = 0; i < 16; i++) fin = fin && isfinite(Tn[i]);
        if (fin) for (int i = 0; i < 16; i++) U[i] = Tn[i];
    } else if (n > 0.0) {
        double A[36], nb[6], x[6];
        int q = 0;
        for (int p = 0; p < 6; p++) for (int r = p; r < 6; r++) { A[p * 6 + r] = tot[q]; A[r * 6 + p] = tot[q]; q++; }
        for (int p = 0; p < 6; p++) nb[p] = -tot[21 + p];
        if (icp_ldlt6(A, nb, x)) icp_vec6_to_m4(x, U);
    }
    for (int i = 0; i < 16; i++) st->U[i] = U[i];
}

struct IcpWs { void* p = nullptr; size_t bytes = 0; };
IcpWs g_icp_ws[16];

int icp_get_ws(size_t bytes, void** out) {
    int dev = 0;
    B2R_CUDA(cudaGetDevice(&dev));
    IcpWs& w = g_icp_ws[dev & 15];
    if (w.bytes < bytes) {
        if (w.p) { cudaDeviceSynchronize(); cudaFree(w.p); }
        w.p = nullptr; w.bytes = 0;
        B2R_CUDA(cudaMalloc(&w.p, bytes));
        w.bytes = bytes;
    }
    *out = w.p;
    return B2R_OK;
}

inline size_t al(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

struct IcpColor { const double* src_int; const double* tgt_int; const double* tgt_grad; double lambda_geometric; };
static const IcpColor kPointToPoint = {nullptr, nullptr, nullptr, -1.0};     // marker: point-to-point estimation

static int icp_run(const double* src_xyz, int64_t N, const double* tgt_xyz, const double* tgt_nrm, int64_t M,
                   double max_dist, const double* init_host, int max_iter, double rel_fitness,
                   double rel_rmse, double* out_T_host, double* fitness_host, double* rmse_host,
                   int32_t* corr, const IcpColor* color, cudaStream_t stream) {
    if (!(max_dist > 0.0)) { b2r_set_error("[RegistrationICP] Invalid max_correspondence_distance."); return B2R_ERR_BAD_ARG; }
    const bool p2p = color == &kPointToPoint;
    if (p2p) color = nullptr;
    if (!tgt_nrm && !p2p) { b2r_set_error(color ? "ColoredICP requires target pointcloud to have normals." : "[RegistrationICP] TransformationEstimationPointToPlane requires target normals."); return B2R_ERR_FORMAT; }
    if (!src_xyz || !tgt_xyz || !init_host || !out_T_host || !fitness_host || !rmse_host || N < 0 || M < 0 || max_iter < 0 ||
        N > 0x7fffffff || M > 0x7fffffff) {
        b2r_set_error("b2r_icp_point_to_plane: bad argument");
        return B2R_ERR_BAD_ARG;
    }
    if (N == 0 || M == 0) {
        memcpy(out_T_host, init_host, 128);
        *fitness_host = 0.0; *rmse_host = 0.0;
        return 0;
    }
    // The per-device workspace is process-global and a ctypes call releases the GIL (the reference runs one thread per
    // camera): one ICP at a time per device; the lock is held until the results have been read back.
    int cur_dev = 0;
    B2R_CUDA(cudaGetDevice(&cur_dev));
    static std::mutex icp_mu[16];
    std::lock_guard<std::mutex> icp_lock(icp_mu[cur_dev & 15]);
    int nsm = 148;
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, cur_dev);
    unsigned grid = (unsigned)((N + ICP_THREADS - 1) / ICP_THREADS);
    if (grid > (unsigned)nsm * 4) grid = (unsigned)nsm * 4;
    const size_t max_cells = (size_t)8 << 20;
    size_t sort_tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp, (unsigned*)nullptr, (unsigned*)nullptr, (int*)nullptr, (int*)nullptr, (int)M);
    size_t total = 0;
    auto take = [&](size_t b) { size_t o = total; total += al(b); return o; };
    const size_t o_st = take(sizeof(IcpState)), o_pcd = take((size_t)N * 24), o_corr = take((size_t)N * 4);
    const size_t o_cin = take((size_t)M * 4), o_cout = take((size_t)M * 4), o_iin = take((size_t)M * 4), o_iout = take((size_t)M * 4);
    const size_t o_cs = take(max_cells * 4), o_ce = take(max_cells * 4);
    const size_t o_part = take((size_t)grid * ICP_NV * 8), o_tmp = take(sort_tmp);
    void* ws = nullptr;
    int rc = icp_get_ws(total, &ws);
    if (rc != B2R_OK) return rc;
    char* b = (char*)ws;
    IcpState* st = (IcpState*)(b + o_st);
    IcpState h0;
    memset(&h0, 0, sizeof(h0));
    memcpy(h0.T, init_host, 128);
    memcpy(h0.U, init_host, 128);
    B2R_CUDA(cudaMemcpyAsync(st, &h0, sizeof(h0), cudaMemcpyHostToDevice, stream));
    bounds_kernel<<<1, 1024, 0, stream>>>(tgt_xyz, (long long)M, st);
    B2R_CHECK_LAUNCH();
    IcpState h1;
    B2R_CUDA(cudaMemcpyAsync(&h1, st, sizeof(h1), cudaMemcpyDeviceToHost, stream));
    B2R_CUDA(cudaStreamSynchronize(stream));
    IcpGrid g;
    g.cell = max_dist;
    for (int k = 0; k < 3; k++) g.lo[k] = h1.bounds[k];
    for (;;) {
        double n = 1.0;
        for (int k = 0; k < 3; k++) { g.dim[k] = (int)fmax(1.0, floor((h1.bounds[3 + k] - g.lo[k]) / g.cell) + 1.0); n *= (double)g.dim[k]; }
        if (n <= (double)max_cells) break;
        g.cell *= 2.0;
    }
    const size_t ncells = (size_t)g.dim[0] * g.dim[1] * g.dim[2];
    unsigned *cin = (unsigned*)(b + o_cin), *cout_ = (unsigned*)(b + o_cout);
    int *iin = (int*)(b + o_iin), *iout = (int*)(b + o_iout);
    int *cs = (int*)(b + o_cs), *ce = (int*)(b + o_ce);
    cell_id_kernel<<<(unsigned)((M + 255) / 256), 256, 0, stream>>>(tgt_xyz, (int)M, g, cin, iin);
    B2R_CHECK_LAUNCH();
    size_t tmpb = sort_tmp;
    B2R_CUDA(cub::DeviceRadixSort::SortPairs(b + o_tmp, tmpb, cin, cout_, iin, iout, (int)M, 0, 32, stream));
    B2R_CUDA(cudaMemsetAsync(cs, 0, ncells * 4, stream));
    B2R_CUDA(cudaMemsetAsync(ce, 0, ncells * 4, stream));
    cell_range_kernel<<<(unsigned)((M + 255) / 256), 256, 0, stream>>>(cout_, (int)M, cs, ce);
    copy_points_kernel<<<(unsigned)((N * 3 + 255) / 256), 256, 0, stream>>>(src_xyz, (double*)(b + o_pcd), (long long)N * 3);
    B2R_CHECK_LAUNCH();
    IcpArgs a;
    a.pcd = (double*)(b + o_pcd); a.N = N; a.tgt = tgt_xyz; a.nrm = tgt_nrm; a.sorted_idx = iout; a.cell_start = cs; a.cell_end = ce;
    a.g = g; a.max_dist = max_dist; a.corr = (int32_t*)(b + o_corr); a.partials = (double*)(b + o_part); a.st = st;
    a.max_iter = max_iter; a.rel_fitness = rel_fitness; a.rel_rmse = rel_rmse;
    a.colored = color ? 1 : 0;
    a.p2p = p2p ? 1 : 0;
    a.src_int = color ? color->src_int : nullptr; a.tgt_int = color ? color->tgt_int : nullptr; a.tgt_grad = color ? color->tgt_grad : nullptr;
    a.sqrt_lg = color ? sqrt(color->lambda_geometric) : 1.0; a.sqrt_lp = color ? sqrt(1.0 - color->lambda_geometric) : 0.0;
    for (int k = 0; k <= max_iter; k++) {
        icp_iter_kernel<<<grid, ICP_THREADS, 0, stream>>>(a);
        if ((k & 15) == 15) {     // cheap early exit: look at the done flag every 16 launches
            B2R_CUDA(cudaMemcpyAsync(&h1, st, sizeof(h1), cudaMemcpyDeviceToHost, stream));
            B2R_CUDA(cudaStreamSynchronize(stream));
            if (h1.done) break;
        }
    }
    B2R_CHECK_LAUNCH();
    B2R_CUDA(cudaMemcpyAsync(&h1, st, sizeof(h1), cudaMemcpyDeviceToHost, stream));
    if (corr) B2R_CUDA(cudaMemcpyAsync(corr, a.corr, (size_t)N * 4, cudaMemcpyDeviceToDevice, stream));
    B2R_CUDA(cudaStreamSynchronize(stream));
    memcpy(out_T_host, h1.T, 128);
    *fitness_host = h1.fitness;
    *rmse_host = h1.rmse;
    return h1.iters;
}


// registration_icp with TransformationEstimationPointToPlane.  See include/b2recon.h.
B2R_API int b2r_icp_point_to_plane(const double* src_xyz, int64_t N, const double* tgt_xyz, const double* tgt_nrm, int64_t M,
                                   double max_dist, const double* init_host, int max_iter, double rel_fitness,
                                   double rel_rmse, double* out_T_host, double* fitness_host, double* rmse_host,
                                   int32_t* corr, cudaStream_t stream) {
    return icp_run(src_xyz, N, tgt_xyz, tgt_nrm, M, max_dist, init_host, max_iter, rel_fitness, rel_rmse, out_T_host, fitness_host,
                   rmse_host, corr, nullptr, stream);
}

// registration_icp with TransformationEstimationPointToPoint(with_scaling=False)
// (CodeThatWorks/Works2CamerasCOLORVisualizer.py:252-257, CodeThatKindaWorks/BetterVisualizerD2cameras.py:248-253).
B2R_API int b2r_icp_point_to_point(const double* src_xyz, int64_t N, const double* tgt_xyz, int64_t M, double max_dist,
                                   const double* init_host, int max_iter, double rel_fitness, double rel_rmse, double* out_T_host,
                                   double* fitness_host, double* rmse_host, int32_t* corr, cudaStream_t stream) {
    return icp_run(src_xyz, N, tgt_xyz, nullptr, M, max_dist, init_host, max_iter, rel_fitness, rel_rmse, out_T_host, fitness_host,
                   rmse_host, corr, &kPointToPoint, stream);
}

// ------------------------------------------------------------------------------------------------
// f4: registration_colored_icp (GoodColorReconstruction.py:246-254; Park et al. 2017 as in Open3D's ColoredICP.cpp)
// ------------------------------------------------------------------------------------------------
struct b2r_knn;
extern "C" b2r_knn* b2r_knn_create(const double* xyz, int64_t n, double cell, int k_hint, cudaStream_t stream);
extern "C" void b2r_knn_destroy(b2r_knn* h);
extern "C" int b2r_knn_search(const b2r_knn* h, const double* query, int64_t nq, int k, double radius, int32_t* idx, double* d2,
                              int32_t* counts, cudaStream_t stream);

namespace {

__global__ void intensity_kernel(const double* __restrict__ rgb, long long n, double* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = xddiv(xdadd(xdadd(rgb[3 * i], rgb[3 * i + 1]), rgb[3 * i + 2]), 3.0);
}

// InitializePointCloudForColoredICP: least-squares colour gradient on the tangent plane of every target point from its
// hybrid-search neighbours (itself first), plus the row (nn-1) * n^T g = 0; normal equations accumulated in row order,
// solved by an unpivoted LDL^T (zero gradient when fewer than 4 neighbours or a pivot is not positive).
__global__ void color_gradient_kernel(const double* __restrict__ xyz, const double* __restrict__ nrm, const double* __restrict__ inten,
                                      const int32_t* __restrict__ nbr, const int32_t* __restrict__ counts, int k, long long M,
                                      double* __restrict__ grad) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    double g[3] = {0.0, 0.0, 0.0};
    const int nn = counts[i];
    if (nn >= 4) {
        const double vt[3] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]}, nt[3] = {nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]};
        const double it = inten[i];
        double S[6] = {0, 0, 0, 0, 0, 0}, rhs[3] = {0, 0, 0};   // xx xy xz yy yz zz
        for (int q = 1; q < nn; q++) {
            const int j = nbr[i * (long long)k + q];
            const double d[3] = {xdsub(xyz[3 * j], vt[0]), xdsub(xyz[3 * j + 1], vt[1]), xdsub(xyz[3 * j + 2], vt[2])};
            const double dn = xdadd(xdadd(xdmul(d[0], nt[0]), xdmul(d[1], nt[1])), xdmul(d[2], nt[2]));
            double a[3];
#pragma unroll
            for (int c = 0; c < 3; c++) a[c] = xdsub(xdsub(xyz[3 * j + c], xdmul(dn, nt[c])), vt[c]);   // vt_proj - vt
            const double b = xdsub(inten[j], it);
            S[0] = xdadd(S[0], xdmul(a[0], a[0])); S[1] = xdadd(S[1], xdmul(a[0], a[1])); S[2] = xdadd(S[2], xdmul(a[0], a[2]));
            S[3] = xdadd(S[3], xdmul(a[1], a[1])); S[4] = xdadd(S[4], xdmul(a[1], a[2])); S[5] = xdadd(S[5], xdmul(a[2], a[2]));
#pragma unroll
            for (int c = 0; c < 3; c++) rhs[c] = xdadd(rhs[c], xdmul(a[c], b));
        }
        const double w = (double)(nn - 1);
        const double a[3] = {xdmul(w, nt[0]), xdmul(w, nt[1]), xdmul(w, nt[2])};
        S[0] = xdadd(S[0], xdmul(a[0], a[0])); S[1] = xdadd(S[1], xdmul(a[0], a[1])); S[2] = xdadd(S[2], xdmul(a[0], a[2]));
        S[3] = xdadd(S[3], xdmul(a[1], a[1])); S[4] = xdadd(S[4], xdmul(a[1], a[2])); S[5] = xdadd(S[5], xdmul(a[2], a[2]));
        // LDL^T of [[S0 S1 S2][S1 S3 S4][S2 S4 S5]]
        const double d0 = S[0];
        if (d0 > 0.0) {
            const double l10 = xddiv(S[1], d0), l20 = xddiv(S[2], d0);
            const double d1 = xdsub(S[3], xdmul(l10, S[1]));
            if (d1 > 0.0) {
                const double l21 = xddiv(xdsub(S[4], xdmul(l20, S[1])), d1);
                const double d2 = xdsub(xdsub(S[5], xdmul(l20, S[2])), xdmul(xdmul(l21, l21), d1));
                if (d2 > 0.0) {
                    const double y0 = rhs[0], y1 = xdsub(rhs[1], xdmul(l10, y0)), y2 = xdsub(xdsub(rhs[2], xdmul(l20, y0)), xdmul(l21, y1));
                    const double z2 = xddiv(y2, d2), z1 = xdsub(xddiv(y1, d1), xdmul(l21, z2));
                    const double z0 = xdsub(xdsub(xddiv(y0, d0), xdmul(l10, z1)), xdmul(l20, z2));
                    g[0] = z0; g[1] = z1; g[2] = z2;
                }
            }
        }
    }
    grad[3 * i] = g[0]; grad[3 * i + 1] = g[1]; grad[3 * i + 2] = g[2];
}

}  // namespace

// registration_colored_icp(source, target, max_dist, init, TransformationEstimationForColoredICP(lambda_geometric), criteria).
// DEVICE f64 arrays; rgb in [0, 1] as PointCloud.colors holds them.  grad_out (optional, device [M][3]) receives the
// target's colour gradients (parity hook).  Returns the iteration count.
B2R_API int b2r_icp_colored(const double* src_xyz, const double* src_rgb, int64_t N, const double* tgt_xyz, const double* tgt_nrm,
                            const double* tgt_rgb, int64_t M, double max_dist, double lambda_geometric, const double* init_host,
                            int max_iter, double rel_fitness, double rel_rmse, double* out_T_host, double* fitness_host,
                            double* rmse_host, int32_t* corr, double* grad_out, cudaStream_t stream) {
    if (!tgt_nrm) { b2r_set_error("ColoredICP requires target pointcloud to have normals."); return B2R_ERR_FORMAT; }
    if (!tgt_rgb) { b2r_set_error("ColoredICP requires target pointcloud to have colors."); return B2R_ERR_FORMAT; }
    if (!src_rgb) { b2r_set_error("ColoredICP requires source pointcloud to have colors."); return B2R_ERR_FORMAT; }
    if (!(lambda_geometric >= 0.0 && lambda_geometric <= 1.0)) { b2r_set_error("TransformationEstimationForColoredICP::lambda_geometric needs to be in range [0, 1]."); return B2R_ERR_BAD_ARG; }
    if (!(max_dist > 0.0)) { b2r_set_error("[RegistrationICP] Invalid max_correspondence_distance."); return B2R_ERR_BAD_ARG; }
    if (!src_xyz || !tgt_xyz || N < 0 || M < 0 || N > 0x7fffffff || M > 0x7fffffff) { b2r_set_error("b2r_icp_colored: bad argument"); return B2R_ERR_BAD_ARG; }
    if (N == 0 || M == 0)
        return icp_run(src_xyz, N, tgt_xyz, tgt_nrm, M, max_dist, init_host, max_iter, rel_fitness, rel_rmse, out_T_host, fitness_host, rmse_host, corr, nullptr, stream);
    const int K = 30;                                       // KDTreeSearchParamHybrid(max_dist * 2, 30)
    double *sint = nullptr, *tint = nullptr, *grad = nullptr;
    int32_t *nbr = nullptr, *cnt = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&sint, (size_t)N * 8, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&tint, (size_t)M * 8, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&grad, (size_t)M * 24, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&nbr, (size_t)M * K * 4, stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&cnt, (size_t)M * 4, stream);
    int rc = B2R_OK;
    if (e != cudaSuccess) { b2r_set_error("b2r_icp_colored: %s", cudaGetErrorString(e)); rc = B2R_ERR_CUDA; }
    b2r_knn* tree = nullptr;
    if (rc == B2R_OK) {
        tree = b2r_knn_create(tgt_xyz, M, max_dist * 2.0, K, stream);
        if (!tree) rc = B2R_ERR_CUDA;
    }
    if (rc == B2R_OK) rc = b2r_knn_search(tree, nullptr, M, K, max_dist * 2.0, nbr, nullptr, cnt, stream);
    if (rc == B2R_OK) {
        intensity_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(src_rgb, (long long)N, sint);
        intensity_kernel<<<(unsigned)((M + 255) / 256), 256, 0, stream>>>(tgt_rgb, (long long)M, tint);
        color_gradient_kernel<<<(unsigned)((M + 127) / 128), 128, 0, stream>>>(tgt_xyz, tgt_nrm, tint, nbr, cnt, K, (long long)M, grad);
        if (cudaGetLastError() != cudaSuccess) { b2r_set_error("b2r_icp_colored: launch failed"); rc = B2R_ERR_CUDA; }
        if (rc == B2R_OK && grad_out && cudaMemcpyAsync(grad_out, grad, (size_t)M * 24, cudaMemcpyDeviceToDevice, stream) != cudaSuccess) rc = B2R_ERR_CUDA;
    }
    if (rc == B2R_OK) {
        const IcpColor col = {sint, tint, grad, lambda_geometric};
        rc = icp_run(src_xyz, N, tgt_xyz, tgt_nrm, M, max_dist, init_host, max_iter, rel_fitness, rel_rmse, out_T_host, fitness_host, rmse_host, corr, &col, stream);
    }
    if (tree) b2r_knn_destroy(tree);
    cudaFreeAsync(sint, stream); cudaFreeAsync(tint, stream); cudaFreeAsync(grad, stream); cudaFreeAsync(nbr, stream); cudaFreeAsync(cnt, stream);
    return rc;
}
