// oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement (C++17 + OpenMP) of the arithmetic the reference scripts reach through
// Open3D 0.19.0's legacy API (open3d==0.19.0, pinned at /root/reference/pyproject.toml:15).
// None of that arithmetic lives under /root/reference (the scripts are call sites only), so
// every function below cites (a) the reference call site whose contract it serves and (b) the
// upstream semantics as written out in SURVEY.md Appendix A.
//
// PARITY UNPINNED: Open3D cannot be installed or built in this environment and the reference
// holds no tests, golden vectors or fixtures for this path.  The oracle is
// therefore pinned only by the analytic known-answer tests in tests/ (closed-form planes,
// spheres, rigid motions) and by the committed fixtures in tests/golden/.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product library (libb2recon.so) never links or calls it.
//
// Build: see oracle/Makefile  (g++ -O3 -fopenmp -ffp-contract=off; contraction is disabled so
// f32/f64 expressions round exactly as a non-FMA x86-64 Open3D wheel would).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <memory>
#include <unordered_map>
#include <unordered_set>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MC_TABLE_QUAL static const
#include "../azurekinect3dreconstruction_b200/csrc/mc_tables.h"   // the one generated copy of the constant tables (tools/gen_mc_tables.py)

#define O3R_API extern "C" __attribute__((visibility("default")))

namespace {

// ----------------------------------------------------------------------------------------------
// small linear algebra (row-major)
// ----------------------------------------------------------------------------------------------
struct M4 {
    double m[16];
    double operator()(int r, int c) const { return m[r * 4 + c]; }
    double& operator()(int r, int c) { return m[r * 4 + c]; }
};

M4 m4_identity() {
    M4 r{};
    for (int i = 0; i < 4; i++) r(i, i) = 1.0;
    return r;
}

M4 m4_mul(const M4& a, const M4& b) {
    M4 r{};
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            double s = 0.0;
            for (int k = 0; k < 4; k++) s += a(i, k) * b(k, j);
            r(i, j) = s;
        }
    return r;
}

// General 4x4 inverse by cofactor expansion (the closed form Eigen uses for fixed 4x4).  The
// product library carries the same operation sequence (csrc/host_math.h) so that
// camera_pose = extrinsic^-1 is bit-identical on both sides.
bool m4_inverse(const M4& a, M4* out) {
    const double* m = a.m;
    double inv[16];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    if (det == 0.0 || !std::isfinite(det)) return false;
    double idet = 1.0 / det;
    for (int i = 0; i < 16; i++) out->m[i] = inv[i] * idet;
    return true;
}

// p' = (M * (x,y,z,1)).xyz with the sequential column accumulation ((m0*x + m1*y) + m2*z) + m3.
inline void m4_apply(const M4& M, double x, double y, double z, double* o) {
    for (int i = 0; i < 3; i++) o[i] = ((M(i, 0) * x + M(i, 1) * y) + M(i, 2) * z) + M(i, 3);
}

// 6x6 symmetric solve by LDLT with diagonal pivoting (Eigen::LDLT's strategy), SURVEY A.6.
bool ldlt6_solve(const double A_in[36], const double b_in[6], double x[6]) {
    const int n = 6;
    double A[36];
    std::memcpy(A, A_in, sizeof(A));
    int perm[6];
    for (int i = 0; i < n; i++) perm[i] = i;
    double D[6];
    double L[36] = {0};
    // work on a permuted copy, symmetric full storage
    for (int k = 0; k < n; k++) {
        // pivot: largest |diagonal| of the trailing block
        int p = k;
        double best = std::fabs(A[k * n + k]);
        for (int i = k + 1; i < n; i++)
            if (std::fabs(A[i * n + i]) > best) { best = std::fabs(A[i * n + i]); p = i; }
        if (p != k) {
            for (int j = 0; j < n; j++) std::swap(A[k * n + j], A[p * n + j]);
            for (int j = 0; j < n; j++) std::swap(A[j * n + k], A[j * n + p]);
            for (int j = 0; j < k; j++) std::swap(L[k * n + j], L[p * n + j]);
            std::swap(perm[k], perm[p]);
        }
        double d = A[k * n + k];
        D[k] = d;
        L[k * n + k] = 1.0;
        if (d == 0.0) {  // zero pivot: remaining block is (numerically) zero; Eigen stops here
            for (int i = k + 1; i < n; i++) { D[i] = 0.0; L[i * n + i] = 1.0; }
            for (int i = k + 1; i < n; i++) L[i * n + k] = 0.0;
            // finish permutation bookkeeping and break
            for (int kk = k + 1; kk < n; kk++) D[kk] = 0.0;
            break;
        }
        for (int i = k + 1; i < n; i++) L[i * n + k] = A[i * n + k] / d;
        for (int i = k + 1; i < n; i++)
            for (int j = k + 1; j < n; j++) A[i * n + j] -= L[i * n + k] * d * L[j * n + k];
    }
    // solve  P^T L D L^T P x = b
    double y[6], z[6];
    for (int i = 0; i < n; i++) y[i] = b_in[perm[i]];
    for (int i = 0; i < n; i++) {
        double s = y[i];
        for (int j = 0; j < i; j++) s -= L[i * n + j] * z[j];
        z[i] = s;
    }
    double dmax = 0.0;
    for (int i = 0; i < n; i++) dmax = std::max(dmax, std::fabs(D[i]));
    const double tol = std::numeric_limits<double>::min();
    for (int i = 0; i < n; i++) z[i] = (std::fabs(D[i]) > tol) ? z[i] / D[i] : 0.0;
    (void)dmax;
    for (int i = n - 1; i >= 0; i--) {
        double s = z[i];
        for (int j = i + 1; j < n; j++) s -= L[j * n + i] * y[j];
        y[i] = s;
    }
    for (int i = 0; i < n; i++) x[perm[i]] = y[i];
    for (int i = 0; i < n; i++)
        if (!std::isfinite(x[i])) return false;
    return true;
}

// TransformVector6dToMatrix4d: [Rz(x2) Ry(x1) Rx(x0) | x3..5]  
M4 vec6_to_m4(const double x[6]) {
    double ca = std::cos(x[0]), sa = std::sin(x[0]);
    double cb = std::cos(x[1]), sb = std::sin(x[1]);
    double cg = std::cos(x[2]), sg = std::sin(x[2]);
    M4 T = m4_identity();
    T(0, 0) = cg * cb; T(0, 1) = cg * sb * sa - sg * ca; T(0, 2) = cg * sb * ca + sg * sa;
    T(1, 0) = sg * cb; T(1, 1) = sg * sb * sa + cg * ca; T(1, 2) = sg * sb * ca - cg * sa;
    T(2, 0) = -sb;     T(2, 1) = cb * sa;                T(2, 2) = cb * ca;
    T(0, 3) = x[3]; T(1, 3) = x[4]; T(2, 3) = x[5];
    return T;
}

struct NormalEq {
    double JTJ[36];
    double JTr[6];
    double r2;
    void zero() { std::memset(this, 0, sizeof(*this)); }
    inline void add_row(const double J[6], double r) {
        for (int a = 0; a < 6; a++) {
            for (int b = 0; b < 6; b++) JTJ[a * 6 + b] += J[a] * J[b];
            JTr[a] += J[a] * r;
        }
        r2 += r * r;
    }
    void merge(const NormalEq& o) {
        for (int i = 0; i < 36; i++) JTJ[i] += o.JTJ[i];
        for (int i = 0; i < 6; i++) JTr[i] += o.JTr[i];
        r2 += o.r2;
    }
};

bool solve_update(const NormalEq& ne, M4* dT) {
    double nb[6], x[6];
    for (int i = 0; i < 6; i++) nb[i] = -ne.JTr[i];
    if (!ldlt6_solve(ne.JTJ, nb, x)) { *dT = m4_identity(); return false; }
    *dT = vec6_to_m4(x);
    return true;
}

}  // namespace

// ================================================================================================
// a1  RGBDImage.create_from_color_and_depth  (reference: CodeThatKindaWorks/Depth2PC.py:60-66,
//     CodeThatWorks/GoodColorReconstruction.py:131-137; upstream semantics SURVEY A.1)
// ================================================================================================
O3R_API void o3r_depth_u16_to_f32(const uint16_t* depth, int64_t n, double depth_scale, double depth_trunc, float* out) {
    const float scale = (float)depth_scale, trunc = (float)depth_trunc;
    for (int64_t i = 0; i < n; i++) {
        float p = (float)depth[i];
        p /= scale;
        if (p >= trunc) p = 0.0f;
        out[i] = p;
    }
}

// 3-channel u8 -> f32 intensity (CreateFloatImage, weighted), RGBD_Odometry.py:15-18 default path
O3R_API void o3r_rgb_to_intensity(const uint8_t* rgb, int64_t n, float* out) {
    for (int64_t i = 0; i < n; i++) {
        const uint8_t* p = rgb + 3 * i;
        out[i] = (0.2990f * (float)p[0] + 0.5870f * (float)p[1] + 0.1140f * (float)p[2]) / 255.0f;
    }
}

// ================================================================================================
// a3 / a4  PointCloud.create_from_rgbd_image / create_from_depth_image
//     (reference: Depth2PC.py:69-84, BetterVisualizer2CamWorksCOLOR.py:528; upstream SURVEY A.3)
//     colour: rgb (u8 x3) or gray (f32 x1) or neither.  Returns the number of points written.
//     xyz/col may be null for a pure count.  src_idx receives the raster index i*W+j per point.
// ================================================================================================
O3R_API int64_t o3r_unproject(const float* depth, int W, int H, const uint8_t* rgb, const float* gray,
                              double fx, double fy, double cx, double cy, const double* extrinsic, int stride,
                              int valid_only, double* xyz, double* col, int32_t* src_idx) {
    M4 E, pose;
    std::memcpy(E.m, extrinsic, sizeof(E.m));
    if (!m4_inverse(E, &pose)) return -1;
    int64_t cnt = 0;
    const double nan = std::numeric_limits<double>::quiet_NaN();
    for (int i = 0; i < H; i += stride)
        for (int j = 0; j < W; j += stride) {
            float d = depth[(int64_t)i * W + j];
            bool ok = d > 0;
            if (!ok && valid_only) continue;
            if (xyz) {
                if (ok) {
                    double z = (double)d;
                    double x = (j - cx) * z / fx;
                    double y = (i - cy) * z / fy;
                    m4_apply(pose, x, y, z, xyz + 3 * cnt);
                } else {
                    xyz[3 * cnt] = xyz[3 * cnt + 1] = xyz[3 * cnt + 2] = nan;
                }
            }
            if (col) {
                if (rgb) {
                    const uint8_t* p = rgb + 3 * ((int64_t)i * W + j);
                    col[3 * cnt] = p[0] / 255.0; col[3 * cnt + 1] = p[1] / 255.0; col[3 * cnt + 2] = p[2] / 255.0;
                } else if (gray) {
                    double g = (double)gray[(int64_t)i * W + j];
                    col[3 * cnt] = col[3 * cnt + 1] = col[3 * cnt + 2] = g;
                }
            }
            if (src_idx) src_idx[cnt] = i * W + j;
            cnt++;
        }
    return cnt;
}

// a5  pcd.transform(T)  (Depth2PC.py:87-93, BetterVisualizer2CamWorksCOLOR.py:634-640)
O3R_API void o3r_transform_points(double* xyz, int64_t n, const double* T, double* normals) {
    for (int64_t k = 0; k < n; k++) {
        double* p = xyz + 3 * k;
        double h[4];
        for (int i = 0; i < 4; i++) h[i] = ((T[i * 4 + 0] * p[0] + T[i * 4 + 1] * p[1]) + T[i * 4 + 2] * p[2]) + T[i * 4 + 3];
        p[0] = h[0] / h[3]; p[1] = h[1] / h[3]; p[2] = h[2] / h[3];
        if (normals) {
            double* q = normals + 3 * k;
            double a = q[0], b = q[1], c = q[2];
            for (int i = 0; i < 3; i++) q[i] = (T[i * 4 + 0] * a + T[i * 4 + 1] * b) + T[i * 4 + 2] * c;
        }
    }
}

// a6  colour-to-depth gather (north_star subsystem 1; replaces libk4a's transformation, which the
//     reference reaches via capture.transformed_depth, e.g. GoodColorReconstruction.py:346-352).
//     libk4a is closed source, so the arithmetic below is this project's own DEFINITION of the nearest
//     colour texel of a depth pixel (unproject with the depth K, rigid depth->colour
//     transform, project with the colour K, nearest texel = floor(u + 0.5)); it is stated once, here, as
//     an f32 chain of correctly rounded fused multiply-adds over constants folded in f64, so that a CPU
//     (std::fmaf) and a GPU (FFMA) evaluate it bit-identically:
//         ax = f32(1/fx)  bx = f32(-cx/fx)  ay = f32(1/fy)  by = f32(-cy/fy)           (normalised ray of a pixel)
//         M  = f32( [[Fx,0,Cx+0.5],[0,Fy,Cy+0.5],[0,0,1]] * [R|t] )                    (3x4, rows u, v, z)
//         xn = fma(u, ax, bx)   yn = fma(v, ay, by)
//         d_r = fma(M[r][0], xn, fma(M[r][1], yn, M[r][2]))     (U,V,Z)_r = fma(z, d_r, M[r][3])
//         Z > 0:  inv = RN(1/Z);  uf = RN(U*inv);  vf = RN(V*inv);  inside iff +0 <= uf < Wc and +0 <= vf < Hc
//     Returns per depth pixel the colour texel index vc*Wc+uc or -1.
O3R_API void o3r_gather_constants(double dfx, double dfy, double dcx, double dcy, double cfx, double cfy, double ccx,
                                  double ccy, const double* d2c, float* out16) {
    out16[0] = (float)(1.0 / dfx); out16[1] = (float)(-dcx / dfx);
    out16[2] = (float)(1.0 / dfy); out16[3] = (float)(-dcy / dfy);
    const double cu = ccx + 0.5, cv = ccy + 0.5;
    for (int j = 0; j < 4; j++) {
        out16[4 + j] = (float)(cfx * d2c[j] + cu * d2c[8 + j]);
        out16[8 + j] = (float)(cfy * d2c[4 + j] + cv * d2c[8 + j]);
        out16[12 + j] = (float)d2c[8 + j];
    }
}

O3R_API void o3r_color_gather_index(const uint16_t* depth, int W, int H, double depth_scale, double depth_trunc,
                                    double dfx, double dfy, double dcx, double dcy, int Wc, int Hc, double cfx,
                                    double cfy, double ccx, double ccy, const double* d2c, int32_t* out_index) {
    float c[16];
    o3r_gather_constants(dfx, dfy, dcx, dcy, cfx, cfy, ccx, ccy, d2c, c);
    const float* M = c + 4;
    const float scale = (float)depth_scale, trunc = (float)depth_trunc;
    const float fW = (float)Wc, fH = (float)Hc;
    for (int v = 0; v < H; v++)
        for (int u = 0; u < W; u++) {
            float z = (float)depth[(int64_t)v * W + u];
            z /= scale;
            if (z >= trunc) z = 0.0f;
            int32_t idx = -1;
            if (z > 0.0f) {
                const float xn = std::fmaf((float)u, c[0], c[1]);
                const float yn = std::fmaf((float)v, c[2], c[3]);
                const float du = std::fmaf(M[0], xn, std::fmaf(M[1], yn, M[2]));
                const float dv = std::fmaf(M[4], xn, std::fmaf(M[5], yn, M[6]));
                const float dz = std::fmaf(M[8], xn, std::fmaf(M[9], yn, M[10]));
                const float U = std::fmaf(z, du, M[3]), V = std::fmaf(z, dv, M[7]), Z = std::fmaf(z, dz, M[11]);
                if (Z > 0.0f) {
                    const float inv = 1.0f / Z;
                    const float uf = U * inv, vf = V * inv;
                    // inside iff the sign bit is clear and the value is below the size (NaN and -0.0 are outside)
                    if (!std::signbit(uf) && uf < fW && !std::signbit(vf) && vf < fH) idx = (int32_t)vf * Wc + (int32_t)uf;
                }
            }
            out_index[(int64_t)v * W + u] = idx;
        }
}

// ================================================================================================
// a7..a11  ScalableTSDFVolume  (reference: GoodColorReconstruction.py:111-115,270,285,305;
//          bundlefusion_reconstruction.py:108-112,276,309-320; upstream SURVEY A.4 / A.5)
// ================================================================================================
namespace {

struct Voxel {
    float tsdf = 0.f;
    float weight = 0.f;
    double color[3] = {0, 0, 0};
};

struct Key {
    int x, y, z;
    bool operator==(const Key& o) const { return x == o.x && y == o.y && z == o.z; }
};
struct KeyHash {
    size_t operator()(const Key& k) const {
        uint64_t h = (uint64_t)(uint32_t)k.x * 0x9E3779B97F4A7C15ull;
        h ^= (uint64_t)(uint32_t)k.y * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
        h ^= (uint64_t)(uint32_t)k.z * 0x165667B19E3779F9ull + (h << 6) + (h >> 2);
        return (size_t)h;
    }
};

struct Unit {
    Key index;
    double origin[3];
    std::vector<Voxel> voxels;
};

struct Volume {
    double voxel_length, sdf_trunc, unit_length;
    int color_type;  // 0 NoColor, 1 RGB8, 2 Gray32
    int res, stride;
    int parallel_units = 0;  // 0: upstream structure (omp over x inside a unit); 1: omp over units
    std::unordered_map<Key, int, KeyHash> map;
    std::vector<std::unique_ptr<Unit>> units;
    int64_t last_updated = 0;  // voxels whose record changed in the last integrate (N_upd)
    int64_t last_touched = 0;
    int64_t total_updated = 0;  // the same, summed over every integrate since the last reset
    int64_t total_touched = 0;  // (unit, frame) pairs since the last reset

    int find(const Key& k) const {
        auto it = map.find(k);
        return it == map.end() ? -1 : it->second;
    }
    int open(const Key& k) {
        auto it = map.find(k);
        if (it != map.end()) return it->second;
        auto u = std::make_unique<Unit>();
        u->index = k;
        u->origin[0] = (double)k.x * unit_length;
        u->origin[1] = (double)k.y * unit_length;
        u->origin[2] = (double)k.z * unit_length;
        u->voxels.resize((size_t)res * res * res);
        int id = (int)units.size();
        units.push_back(std::move(u));
        map.emplace(k, id);
        return id;
    }
    inline int locate(double p) const { return (int)std::floor(p / unit_length); }

    // voxel lookup across unit borders (SURVEY A.5 "V(unit_index, idx)")
    inline const Voxel* lookup(Key unit, int x, int y, int z) const {
        if (x >= res) { x -= res; unit.x += 1; }
        if (y >= res) { y -= res; unit.y += 1; }
        if (z >= res) { z -= res; unit.z += 1; }
        int id = find(unit);
        if (id < 0) return nullptr;
        return &units[id]->voxels[((size_t)x * res + y) * res + z];
    }
};

// UniformTSDFVolume::IntegrateWithDepthToCameraDistanceMultiplier, one unit.
int64_t integrate_unit(Volume& V, Unit& U, const float* depth, const uint8_t* rgb, const float* gray, int W, int H,
                       double dfx, double dfy, double dcx, double dcy, const M4& extrinsic, const float* mult,
                       bool omp_x) {
    const float fx = (float)dfx, fy = (float)dfy, cx = (float)dcx, cy = (float)dcy;
    float E[16];
    for (int i = 0; i < 16; i++) E[i] = (float)extrinsic.m[i];
    const float vl = (float)V.voxel_length;
    const float hv = vl * 0.5f;
    const float trunc = (float)V.sdf_trunc;
    const float inv_trunc = 1.0f / trunc;
    const float Es02 = E[2] * vl, Es12 = E[6] * vl, Es22 = E[10] * vl;
    const float safe_w = (float)W - 0.0001f, safe_h = (float)H - 0.0001f;
    const int res = V.res;
    int64_t updated = 0;
#pragma omp parallel for schedule(static) reduction(+ : updated) if (omp_x)
    for (int x = 0; x < res; x++) {
        for (int y = 0; y < res; y++) {
            const float px = (float)((double)(hv + vl * (float)x) + U.origin[0]);
            const float py = (float)((double)(hv + vl * (float)y) + U.origin[1]);
            const float pz = (float)((double)hv + U.origin[2]);
            float pcx = ((E[0] * px + E[1] * py) + E[2] * pz) + E[3];
            float pcy = ((E[4] * px + E[5] * py) + E[6] * pz) + E[7];
            float pcz = ((E[8] * px + E[9] * py) + E[10] * pz) + E[11];
            Voxel* col = &U.voxels[((size_t)x * res + y) * res];
            for (int z = 0; z < res; z++, pcx += Es02, pcy += Es12, pcz += Es22) {
                if (pcz <= 0) continue;
                float u_f = pcx * fx / pcz + cx + 0.5f;
                float v_f = pcy * fy / pcz + cy + 0.5f;
                if (!(u_f >= 0.0001f && u_f < safe_w && v_f >= 0.0001f && v_f < safe_h)) continue;
                int u = (int)u_f, v = (int)v_f;
                float d = depth[(int64_t)v * W + u];
                if (d <= 0.0f) continue;
                float sdf = (d - pcz) * mult[(int64_t)v * W + u];
                if (sdf > -trunc) {
                    float t = std::min(1.0f, sdf * inv_trunc);
                    Voxel& vx = col[z];
                    vx.tsdf = (vx.tsdf * vx.weight + t) / (vx.weight + 1.0f);
                    if (V.color_type == 1) {
                        const uint8_t* c = rgb + 3 * ((int64_t)v * W + u);
                        const double w = (double)vx.weight, w1 = (double)(vx.weight + 1.0f);
                        for (int k = 0; k < 3; k++) vx.color[k] = (vx.color[k] * w + (double)c[k]) / w1;
                    } else if (V.color_type == 2) {
                        const double g = (double)gray[(int64_t)v * W + u];
                        const double w = (double)vx.weight, w1 = (double)(vx.weight + 1.0f);
                        for (int k = 0; k < 3; k++) vx.color[k] = (vx.color[k] * w + g) / w1;
                    }
                    vx.weight += 1.0f;
                    updated++;
                }
            }
        }
    }
    return updated;
}

}  // namespace

O3R_API void* o3r_tsdf_create(double voxel_length, double sdf_trunc, int color_type, int res, int stride) {
    auto* v = new Volume();
    v->voxel_length = voxel_length;
    v->sdf_trunc = sdf_trunc;
    v->color_type = color_type;
    v->res = res;
    v->stride = stride;
    v->unit_length = voxel_length * res;
    return v;
}
O3R_API void o3r_tsdf_destroy(void* h) { delete (Volume*)h; }
O3R_API void o3r_tsdf_reset(void* h) {
    auto* v = (Volume*)h;
    v->map.clear();
    v->units.clear();
    v->last_updated = v->last_touched = v->total_updated = v->total_touched = 0;
}
O3R_API void o3r_tsdf_set_parallel_units(void* h, int on) { ((Volume*)h)->parallel_units = on; }
O3R_API int64_t o3r_tsdf_num_units(void* h) { return (int64_t)((Volume*)h)->units.size(); }
O3R_API int64_t o3r_tsdf_last_updated(void* h) { return ((Volume*)h)->last_updated; }
O3R_API int64_t o3r_tsdf_last_touched(void* h) { return ((Volume*)h)->last_touched; }
O3R_API int64_t o3r_tsdf_total_updated(void* h) { return ((Volume*)h)->total_updated; }
O3R_API int64_t o3r_tsdf_total_touched(void* h) { return ((Volume*)h)->total_touched; }

// ScalableTSDFVolume::Integrate.  depth: f32 metres W x H; colour: u8x3 (RGB8) or f32 (Gray32).
// Returns 0, or -1 on the upstream "Unsupported image format" condition (checked by the caller's
// shim from shapes; here only null-pointer consistency), -2 singular extrinsic.
O3R_API int o3r_tsdf_integrate(void* h, const float* depth, const uint8_t* rgb, const float* gray, int W, int H,
                               double fx, double fy, double cx, double cy, const double* extrinsic) {
    auto& V = *(Volume*)h;
    if (V.color_type == 1 && !rgb) return -1;
    if (V.color_type == 2 && !gray) return -1;
    M4 E, pose;
    std::memcpy(E.m, extrinsic, sizeof(E.m));
    if (!m4_inverse(E, &pose)) return -2;
    // CreateDepthToCameraDistanceMultiplierFloatImage
    std::vector<float> mult((size_t)W * H), xx(W), yy(H);
    {
        const float ifx = 1.0f / (float)fx, ify = 1.0f / (float)fy, fcx = (float)cx, fcy = (float)cy;
        for (int j = 0; j < W; j++) xx[j] = ((float)j - fcx) * ifx;
        for (int i = 0; i < H; i++) yy[i] = ((float)i - fcy) * ify;
#pragma omp parallel for schedule(static)
        for (int i = 0; i < H; i++)
            for (int j = 0; j < W; j++) mult[(size_t)i * W + j] = std::sqrt(xx[j] * xx[j] + yy[i] * yy[i] + 1.0f);
    }
    // stride-sampled world points -> touched units
    std::unordered_set<Key, KeyHash> touched;
    std::vector<int> order;  // unit ids in first-touch order
    const double tr = V.sdf_trunc;
    for (int i = 0; i < H; i += V.stride)
        for (int j = 0; j < W; j += V.stride) {
            float d = depth[(int64_t)i * W + j];
            if (!(d > 0)) continue;
            double z = (double)d, x = (j - cx) * z / fx, y = (i - cy) * z / fy, p[3];
            m4_apply(pose, x, y, z, p);
            int lo[3], hi[3];
            for (int k = 0; k < 3; k++) { lo[k] = V.locate(p[k] - tr); hi[k] = V.locate(p[k] + tr); }
            for (int a = lo[0]; a <= hi[0]; a++)
                for (int b = lo[1]; b <= hi[1]; b++)
                    for (int c = lo[2]; c <= hi[2]; c++) {
                        Key k{a, b, c};
                        if (touched.insert(k).second) order.push_back(V.open(k));
                    }
        }
    int64_t updated = 0;
    if (V.parallel_units) {
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : updated)
        for (size_t t = 0; t < order.size(); t++)
            updated += integrate_unit(V, *V.units[order[t]], depth, rgb, gray, W, H, fx, fy, cx, cy, E, mult.data(), false);
    } else {
        for (size_t t = 0; t < order.size(); t++)
            updated += integrate_unit(V, *V.units[order[t]], depth, rgb, gray, W, H, fx, fy, cx, cy, E, mult.data(), true);
    }
    V.last_updated = updated;
    V.last_touched = (int64_t)order.size();
    V.total_updated += updated;
    V.total_touched += (int64_t)order.size();
    return 0;
}

O3R_API void o3r_tsdf_export_keys(void* h, int32_t* keys3) {
    auto& V = *(Volume*)h;
    for (size_t i = 0; i < V.units.size(); i++) {
        keys3[3 * i] = V.units[i]->index.x; keys3[3 * i + 1] = V.units[i]->index.y; keys3[3 * i + 2] = V.units[i]->index.z;
    }
}

// Export one unit (x*res*res + y*res + z order). rgb is 3 planes?  No: interleaved [n][3] f64.
O3R_API int o3r_tsdf_export_unit(void* h, const int32_t* key, float* tsdf, float* weight, double* rgb) {
    auto& V = *(Volume*)h;
    int id = V.find(Key{key[0], key[1], key[2]});
    if (id < 0) return -1;
    const auto& vox = V.units[id]->voxels;
    for (size_t i = 0; i < vox.size(); i++) {
        tsdf[i] = vox[i].tsdf; weight[i] = vox[i].weight;
        if (rgb) { rgb[3 * i] = vox[i].color[0]; rgb[3 * i + 1] = vox[i].color[1]; rgb[3 * i + 2] = vox[i].color[2]; }
    }
    return 0;
}

// Bulk export of units [first, first + count) in the oracle's own unit order (first-touch order), as planes:
// keys3 [count][3], tsdf / weight [count][res^3], rgb [count][3][res^3] f64 (may be null) -- the layout the full-size
// parity tests upload and compare on the GPU in one go.
O3R_API void o3r_tsdf_export_range(void* h, int64_t first, int64_t count, int32_t* keys3, float* tsdf, float* weight, double* rgb) {
    auto& V = *(Volume*)h;
    const size_t nv = (size_t)V.res * V.res * V.res;
#pragma omp parallel for schedule(static)
    for (int64_t u = 0; u < count; u++) {
        const Unit& U = *V.units[(size_t)(first + u)];
        keys3[3 * u] = U.index.x; keys3[3 * u + 1] = U.index.y; keys3[3 * u + 2] = U.index.z;
        for (size_t i = 0; i < nv; i++) {
            tsdf[(size_t)u * nv + i] = U.voxels[i].tsdf;
            weight[(size_t)u * nv + i] = U.voxels[i].weight;
            if (rgb)
                for (int c = 0; c < 3; c++) rgb[((size_t)u * 3 + c) * nv + i] = U.voxels[i].color[c];
        }
    }
}

// Test hook: write one unit's voxels directly (used to build analytic SDF volumes for MC tests).
O3R_API void o3r_tsdf_import_unit(void* h, const int32_t* key, const float* tsdf, const float* weight, const double* rgb) {
    auto& V = *(Volume*)h;
    int id = V.open(Key{key[0], key[1], key[2]});
    auto& vox = V.units[id]->voxels;
    for (size_t i = 0; i < vox.size(); i++) {
        vox[i].tsdf = tsdf[i]; vox[i].weight = weight[i];
        if (rgb) { vox[i].color[0] = rgb[3 * i]; vox[i].color[1] = rgb[3 * i + 1]; vox[i].color[2] = rgb[3 * i + 2]; }
    }
}

namespace {

double tsdf_at(const Volume& V, const double p[3]) {
    const double vl = V.voxel_length, hv = 0.5 * vl;
    double q[3] = {p[0] - hv, p[1] - hv, p[2] - hv};
    Key k{V.locate(q[0]), V.locate(q[1]), V.locate(q[2])};
    if (V.find(k) < 0) return 0.0;
    const int ki[3] = {k.x, k.y, k.z};
    int idx[3];
    double r[3];
    for (int i = 0; i < 3; i++) {
        double g = (q[i] - (double)ki[i] * V.unit_length) / vl;
        int f = (int)std::floor(g);
        f = std::min(std::max(f, 0), V.res - 1);
        idx[i] = f;
        r[i] = g - (double)f;
    }
    float f[8];
    for (int i = 0; i < 8; i++) {
        const Voxel* v = V.lookup(k, idx[0] + mc_shift[i][0], idx[1] + mc_shift[i][1], idx[2] + mc_shift[i][2]);
        f[i] = v ? v->tsdf : 0.0f;
    }
    return (1 - r[0]) * ((1 - r[1]) * ((1 - r[2]) * f[0] + r[2] * f[4]) + r[1] * ((1 - r[2]) * f[3] + r[2] * f[7])) +
           r[0] * ((1 - r[1]) * ((1 - r[2]) * f[1] + r[2] * f[5]) + r[1] * ((1 - r[2]) * f[2] + r[2] * f[6]));
}

void normal_at(const Volume& V, const double p[3], double n[3]) {
    const double half_gap = 0.99 * V.voxel_length;
    for (int i = 0; i < 3; i++) {
        double p0[3] = {p[0], p[1], p[2]}, p1[3] = {p[0], p[1], p[2]};
        p0[i] -= half_gap; p1[i] += half_gap;
        n[i] = tsdf_at(V, p1) - tsdf_at(V, p0);
    }
    double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    if (len > 0) { n[0] /= len; n[1] /= len; n[2] /= len; }
}

struct ExtractBuf {
    std::vector<double> a, b, c;
    std::vector<int32_t> t;
    std::vector<int32_t> ids;  // canonical ids per point / vertex: (X,Y,Z,axis)
};
thread_local ExtractBuf g_buf;

}  // namespace

// ScalableTSDFVolume::ExtractPointCloud.  Two-call protocol: the first call computes
// and returns the count (caching the result); o3r_tsdf_extract_points_copy copies it out.
// ids (optional, 4 int32 per point) = global voxel coordinate of p0 and the edge axis: the
// canonical sort key used by the parity tests.
O3R_API int64_t o3r_tsdf_extract_points(void* h) {
    auto& V = *(Volume*)h;
    auto& B = g_buf;
    B.a.clear(); B.b.clear(); B.c.clear(); B.ids.clear();
    const int res = V.res;
    const double vl = V.voxel_length, hv = 0.5 * vl;
    for (const auto& up : V.units) {
        const Unit& U = *up;
        const Key k0 = U.index;
        for (int x = 0; x < res; x++)
            for (int y = 0; y < res; y++)
                for (int z = 0; z < res; z++) {
                    const Voxel& v0 = U.voxels[((size_t)x * res + y) * res + z];
                    const float w0 = v0.weight, f0 = v0.tsdf;
                    if (!(w0 != 0.0f && f0 < 0.98f && f0 >= -0.98f)) continue;
                    double p0[3] = {hv + vl * x + (double)k0.x * V.unit_length, hv + vl * y + (double)k0.y * V.unit_length,
                                    hv + vl * z + (double)k0.z * V.unit_length};
                    for (int i = 0; i < 3; i++) {
                        int d[3] = {0, 0, 0};
                        d[i] = 1;
                        const Voxel* v1 = V.lookup(k0, x + d[0], y + d[1], z + d[2]);
                        if (!v1) continue;
                        const float w1 = v1->weight, f1 = v1->tsdf;
                        if (w1 != 0.0f && f1 < 0.98f && f1 >= -0.98f && f0 * f1 < 0) {
                            const float r0 = std::fabs(f0), r1 = std::fabs(f1);
                            double p[3] = {p0[0], p0[1], p0[2]};
                            double p1i = p0[i] + vl;
                            p[i] = (p0[i] * r1 + p1i * r0) / (r0 + r1);
                            B.a.insert(B.a.end(), p, p + 3);
                            if (V.color_type != 0) {
                                for (int c = 0; c < 3; c++) {
                                    float c0 = (float)v0.color[c], c1 = (float)v1->color[c];
                                    float cc = (c0 * r1 + c1 * r0) / (r0 + r1) / 255.0f;
                                    B.b.push_back((double)cc);
                                }
                            }
                            double n[3];
                            normal_at(V, p, n);
                            B.c.insert(B.c.end(), n, n + 3);
                            B.ids.push_back(k0.x * res + x); B.ids.push_back(k0.y * res + y);
                            B.ids.push_back(k0.z * res + z); B.ids.push_back(i);
                        }
                    }
                }
    }
    return (int64_t)(B.a.size() / 3);
}
O3R_API void o3r_tsdf_extract_points_copy(double* xyz, double* rgb, double* nrm, int32_t* ids) {
    auto& B = g_buf;
    if (xyz) std::memcpy(xyz, B.a.data(), B.a.size() * sizeof(double));
    if (rgb && !B.b.empty()) std::memcpy(rgb, B.b.data(), B.b.size() * sizeof(double));
    if (nrm) std::memcpy(nrm, B.c.data(), B.c.size() * sizeof(double));
    if (ids) std::memcpy(ids, B.ids.data(), B.ids.size() * sizeof(int32_t));
}

// ScalableTSDFVolume::ExtractTriangleMesh.  Returns nV, writes nF via pointer.
namespace {
struct K4 {
    int x, y, z, w;
    bool operator==(const K4& o) const { return x == o.x && y == o.y && z == o.z && w == o.w; }
};
struct K4Hash {
    size_t operator()(const K4& k) const {
        uint64_t h = 1469598103934665603ull;
        for (int v : {k.x, k.y, k.z, k.w}) { h ^= (uint32_t)v; h *= 1099511628211ull; }
        return (size_t)h;
    }
};
}  // namespace

O3R_API int64_t o3r_tsdf_extract_mesh(void* h, int64_t* nF_out) {
    auto& V = *(Volume*)h;
    auto& B = g_buf;
    B.a.clear(); B.b.clear(); B.c.clear(); B.t.clear(); B.ids.clear();
    const int res = V.res;
    const double vl = V.voxel_length, hv = 0.5 * vl;
    std::unordered_map<K4, int, K4Hash> e2v;
    for (const auto& up : V.units) {
        const Unit& U = *up;
        const Key k0 = U.index;
        for (int x = 0; x < res; x++)
            for (int y = 0; y < res; y++)
                for (int z = 0; z < res; z++) {
                    int cube = 0;
                    float w[8], f[8];
                    double c[8][3];
                    for (int i = 0; i < 8; i++) {
                        const Voxel* v = V.lookup(k0, x + mc_shift[i][0], y + mc_shift[i][1], z + mc_shift[i][2]);
                        if (!v) { w[i] = 0.f; f[i] = 0.f; }
                        else {
                            w[i] = v->weight; f[i] = v->tsdf;
                            for (int q = 0; q < 3; q++) c[i][q] = v->color[q] / 255.0;
                        }
                        if (w[i] == 0.0f) { cube = 0; break; }
                        if (f[i] < 0.0f) cube |= (1 << i);
                    }
                    if (cube == 0 || cube == 255) continue;
                    int e2i[12];
                    for (int i = 0; i < 12; i++) {
                        if (!(mc_edge_table[cube] & (1 << i))) continue;
                        K4 e{k0.x * res + x + mc_edge_shift[i][0], k0.y * res + y + mc_edge_shift[i][1],
                             k0.z * res + z + mc_edge_shift[i][2], mc_edge_shift[i][3]};
                        auto it = e2v.find(e);
                        if (it == e2v.end()) {
                            int id = (int)(B.a.size() / 3);
                            e2v.emplace(e, id);
                            e2i[i] = id;
                            double pt[3] = {hv + vl * e.x, hv + vl * e.y, hv + vl * e.z};
                            const int a = mc_edge_to_vert[i][0], b = mc_edge_to_vert[i][1];
                            double f0 = std::fabs((double)f[a]), f1 = std::fabs((double)f[b]);
                            pt[e.w] += f0 * vl / (f0 + f1);
                            B.a.insert(B.a.end(), pt, pt + 3);
                            if (V.color_type != 0)
                                for (int q = 0; q < 3; q++) B.b.push_back((f1 * c[a][q] + f0 * c[b][q]) / (f0 + f1));
                            B.ids.push_back(e.x); B.ids.push_back(e.y); B.ids.push_back(e.z); B.ids.push_back(e.w);
                        } else {
                            e2i[i] = it->second;
                        }
                    }
                    for (int i = 0; mc_tri_table[cube][i] != -1; i += 3) {
                        B.t.push_back(e2i[(int)mc_tri_table[cube][i]]);
                        B.t.push_back(e2i[(int)mc_tri_table[cube][i + 2]]);
                        B.t.push_back(e2i[(int)mc_tri_table[cube][i + 1]]);
                    }
                }
    }
    *nF_out = (int64_t)(B.t.size() / 3);
    return (int64_t)(B.a.size() / 3);
}
O3R_API void o3r_tsdf_extract_mesh_copy(double* vtx, double* vcol, int32_t* tri, int32_t* ids) {
    auto& B = g_buf;
    if (vtx) std::memcpy(vtx, B.a.data(), B.a.size() * sizeof(double));
    if (vcol && !B.b.empty()) std::memcpy(vcol, B.b.data(), B.b.size() * sizeof(double));
    if (tri) std::memcpy(tri, B.t.data(), B.t.size() * sizeof(int32_t));
    if (ids) std::memcpy(ids, B.ids.data(), B.ids.size() * sizeof(int32_t));
}

// TriangleMesh::ComputeVertexNormals (called right after extract_triangle_mesh, e.g.
// GoodColorReconstruction.py:286; SURVEY A.5 last paragraph).
O3R_API void o3r_mesh_vertex_normals(const double* vtx, int64_t nV, const int32_t* tri, int64_t nF, double* out) {
    // upstream: ComputeTriangleNormals(false) -- the UN-normalised cross products (area weighted) are summed per vertex in
    // triangle order, then NormalizeNormals(): Eigen's normalize() leaves a zero vector alone (z > 0 test), and a NaN
    // result becomes (0, 0, 1).
    std::fill(out, out + 3 * nV, 0.0);
    for (int64_t t = 0; t < nF; t++) {
        const double* a = vtx + 3 * tri[3 * t];
        const double* b = vtx + 3 * tri[3 * t + 1];
        const double* c = vtx + 3 * tri[3 * t + 2];
        double e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
        for (int k = 0; k < 3; k++)
            for (int q = 0; q < 3; q++) out[3 * tri[3 * t + k] + q] += n[q];
    }
    for (int64_t v = 0; v < nV; v++) {
        double* n = out + 3 * v;
        double z = n[0] * n[0] + n[1] * n[1] + n[2] * n[2];
        if (z > 0) { double len = std::sqrt(z); n[0] /= len; n[1] /= len; n[2] /= len; }
        if (std::isnan(n[0])) { n[0] = 0.0; n[1] = 0.0; n[2] = 1.0; }
    }
}

// ================================================================================================
// a12  compute_rgbd_odometry  (reference: CodeThatWorks/RGBD_Odometry.py:22-35,
//      CodeThatKindaWorks/capture_frames.py:148-151, bundlefusion_reconstruction.py:118-129,224-227;
//      upstream SURVEY A.6)
// ================================================================================================
namespace {

struct Img {
    int W = 0, H = 0;
    std::vector<float> d;
    Img() {}
    Img(int w, int h) : W(w), H(h), d((size_t)w * h) {}
    float& at(int u, int v) { return d[(size_t)v * W + u]; }
    float at(int u, int v) const { return d[(size_t)v * W + u]; }
};

// horizontal (along u) 3-tap correlation, f64 accumulate, clamp-to-edge
Img filter_h(const Img& in, const double k[3]) {
    Img out(in.W, in.H);
#pragma omp parallel for schedule(static)
    for (int v = 0; v < in.H; v++)
        for (int u = 0; u < in.W; u++) {
            double t = 0.0;
            for (int i = -1; i <= 1; i++) {
                int us = std::min(std::max(u + i, 0), in.W - 1);
                t += (double)in.at(us, v) * k[i + 1];
            }
            out.at(u, v) = (float)t;
        }
    return out;
}
Img filter_v(const Img& in, const double k[3]) {
    Img out(in.W, in.H);
#pragma omp parallel for schedule(static)
    for (int v = 0; v < in.H; v++)
        for (int u = 0; u < in.W; u++) {
            double t = 0.0;
            for (int i = -1; i <= 1; i++) {
                int vs = std::min(std::max(v + i, 0), in.H - 1);
                t += (double)in.at(u, vs) * k[i + 1];
            }
            out.at(u, v) = (float)t;
        }
    return out;
}
const double G3[3] = {0.25, 0.5, 0.25};
const double S31[3] = {-1.0, 0.0, 1.0};
const double S32[3] = {1.0, 2.0, 1.0};
Img gaussian3(const Img& in) { return filter_v(filter_h(in, G3), G3); }
Img sobel_dx(const Img& in) { return filter_v(filter_h(in, S31), S32); }
Img sobel_dy(const Img& in) { return filter_v(filter_h(in, S32), S31); }

Img downsample(const Img& in) {
    Img out(in.W / 2, in.H / 2);
    for (int y = 0; y < out.H; y++)
        for (int x = 0; x < out.W; x++)
            out.at(x, y) = (in.at(2 * x, 2 * y) + in.at(2 * x + 1, 2 * y) + in.at(2 * x, 2 * y + 1) + in.at(2 * x + 1, 2 * y + 1)) / 4.0f;
    return out;
}

struct Corr { int us, vs, ut, vt; };

struct K3 { double fx, fy, cx, cy; };

void compute_correspondence(const K3& K, const M4& E, const Img& ds, const Img& dt, double depth_diff_max,
                            std::vector<Corr>* out) {
    const int W = ds.W, H = ds.H;
    // KRK^-1 and Kt in f64
    double Km[9] = {K.fx, 0, K.cx, 0, K.fy, K.cy, 0, 0, 1};
    double Ki[9] = {1.0 / K.fx, 0, -K.cx / K.fx, 0, 1.0 / K.fy, -K.cy / K.fy, 0, 0, 1};
    double R[9], t[3];
    for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) R[i * 3 + j] = E(i, j); t[i] = E(i, 3); }
    double KR[9], KRKi[9], Kt[3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) { double s = 0; for (int k = 0; k < 3; k++) s += Km[i * 3 + k] * R[k * 3 + j]; KR[i * 3 + j] = s; }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) { double s = 0; for (int k = 0; k < 3; k++) s += KR[i * 3 + k] * Ki[k * 3 + j]; KRKi[i * 3 + j] = s; }
    for (int i = 0; i < 3; i++) { double s = 0; for (int k = 0; k < 3; k++) s += Km[i * 3 + k] * t[k]; Kt[i] = s; }
    std::vector<int> map((size_t)W * H, -1);
    std::vector<float> zbuf((size_t)W * H, -1.0f);
    for (int vs = 0; vs < H; vs++)
        for (int us = 0; us < W; us++) {
            double d = (double)ds.at(us, vs);
            if (std::isnan(d)) continue;
            double q[3];
            for (int i = 0; i < 3; i++)
                q[i] = (((d * KRKi[i * 3 + 0]) * (double)us + (d * KRKi[i * 3 + 1]) * (double)vs) + (d * KRKi[i * 3 + 2])) + Kt[i];
            double zt = q[2];
            int ut = (int)(q[0] / zt + 0.5), vt = (int)(q[1] / zt + 0.5);
            if (ut >= 0 && ut < W && vt >= 0 && vt < H) {
                double dtv = (double)dt.at(ut, vt);
                if (!std::isnan(dtv) && std::fabs(zt - dtv) <= depth_diff_max) {
                    size_t id = (size_t)vt * W + ut;
                    if (map[id] < 0 || (float)d < zbuf[id]) { map[id] = vs * W + us; zbuf[id] = (float)d; }
                }
            }
        }
    out->clear();
    for (int vt = 0; vt < H; vt++)
        for (int ut = 0; ut < W; ut++) {
            int m = map[(size_t)vt * W + ut];
            if (m >= 0) out->push_back(Corr{m % W, m / W, ut, vt});
        }
}

Img preprocess_depth(const Img& in, double dmin, double dmax) {
    Img out = in;
    for (auto& p : out.d)
        if (p < dmin || p > dmax || p <= 0) p = std::numeric_limits<float>::quiet_NaN();
    return out;
}

}  // namespace

// jac_kind: 0 = colour term, 1 = hybrid term.  Returns 1 on success, 0 on failure.
// out_T: 16 doubles row-major (source -> target); out_info: 36 doubles.
// debug_counts (optional, n_levels*? ints): number of correspondences at the first iteration of each level.
O3R_API int o3r_odometry(const float* src_gray, const float* src_depth, const float* tgt_gray, const float* tgt_depth,
                         int W, int H, double fx, double fy, double cx, double cy, const double* init, int jac_kind,
                         const int* iters, int n_levels, double depth_diff_max, double depth_min, double depth_max,
                         double* out_T, double* out_info, int32_t* debug_counts) {
    auto load = [&](const float* p) { Img im(W, H); std::memcpy(im.d.data(), p, sizeof(float) * (size_t)W * H); return im; };
    Img sg = gaussian3(load(src_gray)), tg = gaussian3(load(tgt_gray));
    Img sd = gaussian3(preprocess_depth(load(src_depth), depth_min, depth_max));
    Img td = gaussian3(preprocess_depth(load(tgt_depth), depth_min, depth_max));
    M4 odo;
    std::memcpy(odo.m, init, sizeof(odo.m));
    bool all_zero = true;
    for (double v : odo.m) if (v != 0.0) all_zero = false;
    if (all_zero) odo = m4_identity();
    K3 K0{fx, fy, cx, cy};
    std::vector<Corr> corr;
    compute_correspondence(K0, odo, sd, td, depth_diff_max, &corr);
    {   // NormalizeIntensity
        double ms = 0.0, mt = 0.0;
        for (const auto& c : corr) { ms += (double)sg.at(c.us, c.vs); mt += (double)tg.at(c.ut, c.vt); }
        ms /= (double)corr.size(); mt /= (double)corr.size();
        const double ss = 0.5 / ms, st = 0.5 / mt;
        for (auto& p : sg.d) p = (float)(ss * (double)p);
        for (auto& p : tg.d) p = (float)(st * (double)p);
    }
    // pyramids
    std::vector<Img> sgp{sg}, sdp{sd}, tgp{tg}, tdp{td};
    std::vector<K3> Kp{K0};
    for (int l = 1; l < n_levels; l++) {
        sgp.push_back(downsample(gaussian3(sgp[l - 1])));
        tgp.push_back(downsample(gaussian3(tgp[l - 1])));
        sdp.push_back(downsample(sdp[l - 1]));
        tdp.push_back(downsample(tdp[l - 1]));
        K3 k = Kp[l - 1];
        Kp.push_back(K3{0.5 * k.fx, 0.5 * k.fy, 0.5 * k.cx, 0.5 * k.cy});
    }
    const double sobel_scale = 0.125, lambda = 0.968;
    const double sI = std::sqrt(1.0 - lambda), sD = std::sqrt(lambda);
    M4 result = odo;
    bool ok_all = true;
    for (int level = n_levels - 1; level >= 0 && ok_all; level--) {
        const Img &Sg = sgp[level], &Sd = sdp[level], &Tg = tgp[level], &Td = tdp[level];
        const K3 K = Kp[level];
        Img Tgdx = sobel_dx(Tg), Tgdy = sobel_dy(Tg), Tddx = sobel_dx(Td), Tddy = sobel_dy(Td);
        const int w = Sg.W, hgt = Sg.H;
        // xyz image of the source depth
        std::vector<float> xyz((size_t)w * hgt * 3);
        {
            const double inv_fx = 1.0 / K.fx, inv_fy = 1.0 / K.fy;
            for (int y = 0; y < hgt; y++)
                for (int x = 0; x < w; x++) {
                    float z = Sd.at(x, y);
                    float* o = &xyz[((size_t)y * w + x) * 3];
                    o[0] = (float)((x - K.cx) * z * inv_fx);
                    o[1] = (float)((y - K.cy) * z * inv_fy);
                    o[2] = z;
                }
        }
        const int n_it = iters[n_levels - 1 - level];
        for (int it = 0; it < n_it; it++) {
            compute_correspondence(K, result, Sd, Td, depth_diff_max, &corr);
            if (it == 0 && debug_counts) debug_counts[level] = (int32_t)corr.size();
            NormalEq total;
            total.zero();
            const double R[9] = {result(0, 0), result(0, 1), result(0, 2), result(1, 0), result(1, 1), result(1, 2),
                                 result(2, 0), result(2, 1), result(2, 2)};
            const double t[3] = {result(0, 3), result(1, 3), result(2, 3)};
#pragma omp parallel
            {
                NormalEq priv;
                priv.zero();
#pragma omp for schedule(static) nowait
                for (int64_t ci = 0; ci < (int64_t)corr.size(); ci++) {
                    const Corr c = corr[ci];
                    double diff_photo = (double)(Tg.at(c.ut, c.vt) - Sg.at(c.us, c.vs));
                    double dIdx = sobel_scale * (double)Tgdx.at(c.ut, c.vt);
                    double dIdy = sobel_scale * (double)Tgdy.at(c.ut, c.vt);
                    const float* ps = &xyz[((size_t)c.vs * w + c.us) * 3];
                    double p[3] = {(double)ps[0], (double)ps[1], (double)ps[2]};
                    double P[3];
                    for (int i = 0; i < 3; i++) P[i] = ((R[i * 3] * p[0] + R[i * 3 + 1] * p[1]) + R[i * 3 + 2] * p[2]) + t[i];
                    double invz = 1.0 / P[2];
                    double c0 = dIdx * K.fx * invz, c1 = dIdy * K.fy * invz;
                    double c2 = -(c0 * P[0] + c1 * P[1]) * invz;
                    double J[6] = {-P[2] * c1 + P[1] * c2, P[2] * c0 - P[0] * c2, -P[1] * c0 + P[0] * c1, c0, c1, c2};
                    if (jac_kind == 0) {
                        priv.add_row(J, diff_photo);
                    } else {
                        double J0[6];
                        for (int i = 0; i < 6; i++) J0[i] = sI * J[i];
                        priv.add_row(J0, sI * diff_photo);
                        double dDdx = sobel_scale * (double)Tddx.at(c.ut, c.vt);
                        double dDdy = sobel_scale * (double)Tddy.at(c.ut, c.vt);
                        if (std::isnan(dDdx)) dDdx = 0;
                        if (std::isnan(dDdy)) dDdy = 0;
                        double diff_geo = (double)Td.at(c.ut, c.vt) - P[2];
                        double d0 = dDdx * K.fx * invz, d1 = dDdy * K.fy * invz;
                        double d2 = -(d0 * P[0] + d1 * P[1]) * invz;
                        double J1[6] = {sD * ((-P[2] * d1 + P[1] * d2) - P[1]), sD * ((P[2] * d0 - P[0] * d2) + P[0]),
                                        sD * (-P[1] * d0 + P[0] * d1), sD * d0, sD * d1, sD * (d2 - 1.0)};
                        priv.add_row(J1, sD * diff_geo);
                    }
                }
#pragma omp critical
                total.merge(priv);
            }
            M4 dT;
            bool ok = solve_update(total, &dT);
            result = m4_mul(dT, result);
            if (!ok) { ok_all = false; break; }
        }
    }
    if (!ok_all) {
        M4 I = m4_identity();
        std::memcpy(out_T, I.m, sizeof(I.m));
        for (int i = 0; i < 36; i++) out_info[i] = (i % 7 == 0) ? 1.0 : 0.0;
        return 0;
    }
    std::memcpy(out_T, result.m, sizeof(result.m));
    // CreateInformationMatrix
    compute_correspondence(K0, result, sd, td, depth_diff_max, &corr);
    double G[36] = {0};
    for (const auto& c : corr) {
        float zf = td.at(c.ut, c.vt);
        double inv_fx = 1.0 / fx, inv_fy = 1.0 / fy;
        double x = (double)(float)((c.ut - cx) * zf * inv_fx), y = (double)(float)((c.vt - cy) * zf * inv_fy), z = (double)zf;
        const double g[3][6] = {{0, z, -y, 1, 0, 0}, {-z, 0, x, 0, 1, 0}, {y, -x, 0, 0, 0, 1}};
        for (int r = 0; r < 3; r++)
            for (int a = 0; a < 6; a++)
                for (int b = 0; b < 6; b++) G[a * 6 + b] += g[r][a] * g[r][b];
    }
    std::memcpy(out_info, G, sizeof(G));
    return 1;
}

// Expose the building blocks for unit tests of the GPU image kernels.
O3R_API void o3r_filter(const float* in, int W, int H, int kind, float* out) {
    Img im(W, H);
    std::memcpy(im.d.data(), in, sizeof(float) * (size_t)W * H);
    Img r = kind == 0 ? gaussian3(im) : kind == 1 ? sobel_dx(im) : sobel_dy(im);
    std::memcpy(out, r.d.data(), sizeof(float) * (size_t)W * H);
}
O3R_API void o3r_downsample(const float* in, int W, int H, float* out) {
    Img im(W, H);
    std::memcpy(im.d.data(), in, sizeof(float) * (size_t)W * H);
    Img r = downsample(im);
    std::memcpy(out, r.d.data(), sizeof(float) * r.d.size());
}
// correspondences at pose E; out = 4 int32 per correspondence (us,vs,ut,vt), target raster order.
O3R_API int64_t o3r_correspondence(const float* ds, const float* dt, int W, int H, double fx, double fy, double cx,
                                   double cy, const double* E, double depth_diff_max, int32_t* out, int64_t cap) {
    Img a(W, H), b(W, H);
    std::memcpy(a.d.data(), ds, sizeof(float) * (size_t)W * H);
    std::memcpy(b.d.data(), dt, sizeof(float) * (size_t)W * H);
    M4 e;
    std::memcpy(e.m, E, sizeof(e.m));
    std::vector<Corr> corr;
    compute_correspondence(K3{fx, fy, cx, cy}, e, a, b, depth_diff_max, &corr);
    int64_t n = std::min<int64_t>(cap, (int64_t)corr.size());
    for (int64_t i = 0; i < n; i++) { out[4 * i] = corr[i].us; out[4 * i + 1] = corr[i].vs; out[4 * i + 2] = corr[i].ut; out[4 * i + 3] = corr[i].vt; }
    return (int64_t)corr.size();
}

// ================================================================================================
// a13  registration_icp, TransformationEstimationPointToPlane
//      (reference: GoodColorReconstruction.py:194-198, BetterVisualizer2CamWorksCOLOR.py:255-259;
//       upstream SURVEY A.7).  Exact 1-NN within max_dist through a uniform grid (ties -> lowest j).
// ================================================================================================
namespace {

struct Grid {
    double cell;
    double lo[3];
    int dim[3];
    std::vector<int> start, order;
    const double* pts;
    int64_t M;
    void build(const double* p, int64_t m, double c) {
        pts = p; M = m; cell = c;
        double hi[3] = {-1e300, -1e300, -1e300};
        lo[0] = lo[1] = lo[2] = 1e300;
        for (int64_t i = 0; i < m; i++)
            for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], p[3 * i + k]); hi[k] = std::max(hi[k], p[3 * i + k]); }
        // bound the grid size
        for (;;) {
            double n = 1;
            for (int k = 0; k < 3; k++) { dim[k] = std::max(1, (int)std::floor((hi[k] - lo[k]) / cell) + 1); n *= dim[k]; }
            if (n <= 64e6) break;
            cell *= 2;
        }
        size_t nc = (size_t)dim[0] * dim[1] * dim[2];
        start.assign(nc + 1, 0);
        std::vector<int> cid(m);
        for (int64_t i = 0; i < m; i++) { cid[i] = cell_of(p + 3 * i); start[cid[i] + 1]++; }
        for (size_t c2 = 0; c2 < nc; c2++) start[c2 + 1] += start[c2];
        order.resize(m);
        std::vector<int> fill(start.begin(), start.end() - 1);
        for (int64_t i = 0; i < m; i++) order[fill[cid[i]]++] = (int)i;
    }
    inline int coord(double v, int k) const { return std::min(std::max((int)std::floor((v - lo[k]) / cell), 0), dim[k] - 1); }
    inline int cell_of(const double* p) const { return (coord(p[0], 0) * dim[1] + coord(p[1], 1)) * dim[2] + coord(p[2], 2); }
    // nearest neighbour with d2 <= r2; returns -1 if none
    int nearest(const double* q, double r, double* d2out) const {
        const double r2 = r * r;
        int c0[3], c1[3];
        for (int k = 0; k < 3; k++) {
            c0[k] = (int)std::floor((q[k] - r - lo[k]) / cell);
            c1[k] = (int)std::floor((q[k] + r - lo[k]) / cell);
            if (c1[k] < 0 || c0[k] >= dim[k]) return -1;
            c0[k] = std::max(c0[k], 0); c1[k] = std::min(c1[k], dim[k] - 1);
        }
        int best = -1;
        double bd = r2;
        for (int a = c0[0]; a <= c1[0]; a++)
            for (int b = c0[1]; b <= c1[1]; b++)
                for (int c = c0[2]; c <= c1[2]; c++) {
                    int ci = (a * dim[1] + b) * dim[2] + c;
                    for (int s = start[ci]; s < start[ci + 1]; s++) {
                        int j = order[s];
                        double dx = pts[3 * j] - q[0], dy = pts[3 * j + 1] - q[1], dz = pts[3 * j + 2] - q[2];
                        double d2 = dx * dx + dy * dy + dz * dz;
                        if (d2 > r2) continue;
                        if (best < 0 || d2 < bd || (d2 == bd && j < best)) { bd = d2; best = j; }
                    }
                }
        *d2out = bd;
        return best;
    }
};

struct IcpEval { double fitness, rmse; int64_t n; };

IcpEval icp_evaluate(const std::vector<double>& src, const Grid& g, double max_dist, std::vector<int32_t>& corr) {
    const int64_t N = (int64_t)src.size() / 3;
    corr.assign(N, -1);
    double err2 = 0.0;
    int64_t n = 0;
#pragma omp parallel for schedule(static) reduction(+ : err2, n)
    for (int64_t i = 0; i < N; i++) {
        double d2;
        int j = g.nearest(&src[3 * i], max_dist, &d2);
        if (j >= 0) { corr[i] = j; err2 += d2; n++; }
    }
    IcpEval e;
    e.n = n;
    e.fitness = N ? (double)n / (double)N : 0.0;
    e.rmse = n ? std::sqrt(err2 / (double)n) : 0.0;
    return e;
}

}  // namespace

// Returns the number of iterations executed (>=0) or <0 on the upstream LogError conditions
// (-1: max_dist <= 0, -2: target has no normals).  corr_out (optional, N int32): target index or -1.
O3R_API int o3r_icp_point_to_plane(const double* src_xyz, int64_t N, const double* tgt_xyz, const double* tgt_nrm,
                                   int64_t M, double max_dist, const double* init, int max_iter, double rel_fitness,
                                   double rel_rmse, double* out_T, double* out_fitness, double* out_rmse,
                                   int32_t* corr_out) {
    if (max_dist <= 0.0) return -1;
    if (!tgt_nrm) return -2;
    M4 T;
    std::memcpy(T.m, init, sizeof(T.m));
    std::vector<double> pcd(src_xyz, src_xyz + 3 * N);
    o3r_transform_points(pcd.data(), N, T.m, nullptr);
    Grid g;
    g.build(tgt_xyz, M, max_dist);
    std::vector<int32_t> corr;
    IcpEval res = icp_evaluate(pcd, g, max_dist, corr);
    int it = 0;
    for (; it < max_iter; it++) {
        M4 U = m4_identity();
        if (res.n > 0) {
            NormalEq ne;
            ne.zero();
            for (int64_t i = 0; i < N; i++) {
                int j = corr[i];
                if (j < 0) continue;
                const double* s = &pcd[3 * i];
                const double* t = tgt_xyz + 3 * j;
                const double* n = tgt_nrm + 3 * j;
                double r = (s[0] - t[0]) * n[0] + (s[1] - t[1]) * n[1] + (s[2] - t[2]) * n[2];
                double J[6] = {s[1] * n[2] - s[2] * n[1], s[2] * n[0] - s[0] * n[2], s[0] * n[1] - s[1] * n[0], n[0], n[1], n[2]};
                ne.add_row(J, r);
            }
            if (!solve_update(ne, &U)) U = m4_identity();
        }
        T = m4_mul(U, T);
        o3r_transform_points(pcd.data(), N, U.m, nullptr);
        IcpEval prev = res;
        res = icp_evaluate(pcd, g, max_dist, corr);
        if (std::fabs(prev.fitness - res.fitness) < rel_fitness && std::fabs(prev.rmse - res.rmse) < rel_rmse) { it++; break; }
    }
    std::memcpy(out_T, T.m, sizeof(T.m));
    *out_fitness = res.fitness;
    *out_rmse = res.rmse;
    if (corr_out) std::memcpy(corr_out, corr.data(), sizeof(int32_t) * N);
    return it;
}

// registration_icp with TransformationEstimationPointToPoint(false) (Works2CamerasCOLORVisualizer.py:252-257): same loop, the
// update is the rigid least-squares fit of the correspondences (upstream: Eigen::umeyama; here Horn's closed form, two-pass).
extern "C" void o3r_horn_rigid(const double* s, const double* t, int n, double* T);

O3R_API int o3r_icp_point_to_point(const double* src_xyz, int64_t N, const double* tgt_xyz, int64_t M, double max_dist, const double* init,
                                   int max_iter, double rel_fitness, double rel_rmse, double* out_T, double* out_fitness, double* out_rmse,
                                   int32_t* corr_out) {
    if (max_dist <= 0.0) return -1;
    M4 T;
    std::memcpy(T.m, init, sizeof(T.m));
    std::vector<double> pcd(src_xyz, src_xyz + 3 * N);
    o3r_transform_points(pcd.data(), N, T.m, nullptr);
    Grid g;
    g.build(tgt_xyz, M, max_dist);
    std::vector<int32_t> corr;
    IcpEval res = icp_evaluate(pcd, g, max_dist, corr);
    int it = 0;
    std::vector<double> ps, pt;
    for (; it < max_iter; it++) {
        M4 U = m4_identity();
        if (res.n > 0) {
            ps.clear(); pt.clear();
            for (int64_t i = 0; i < N; i++) {
                if (corr[i] < 0) continue;
                for (int c = 0; c < 3; c++) { ps.push_back(pcd[3 * i + c]); pt.push_back(tgt_xyz[3 * (int64_t)corr[i] + c]); }
            }
            o3r_horn_rigid(ps.data(), pt.data(), (int)(ps.size() / 3), U.m);
            for (int k = 0; k < 16; k++) if (!std::isfinite(U.m[k])) { U = m4_identity(); break; }
        }
        T = m4_mul(U, T);
        o3r_transform_points(pcd.data(), N, U.m, nullptr);
        IcpEval prev = res;
        res = icp_evaluate(pcd, g, max_dist, corr);
        if (std::fabs(prev.fitness - res.fitness) < rel_fitness && std::fabs(prev.rmse - res.rmse) < rel_rmse) { it++; break; }
    }
    std::memcpy(out_T, T.m, sizeof(T.m));
    *out_fitness = res.fitness;
    *out_rmse = res.rmse;
    if (corr_out) std::memcpy(corr_out, corr.data(), sizeof(int32_t) * N);
    return it;
}

// ---- f4: registration_colored_icp (GoodColorReconstruction.py:246-254) ------------------------------------------------
// Open3D 0.19 pipelines/registration/ColoredICP.cpp: InitializePointCloudForColoredICP (colour gradient per target point from
// its SearchHybrid(2 * max_dist, 30) neighbours, least squares on the tangent plane + orthogonality row) and
// TransformationEstimationForColoredICP::ComputeTransformation (geometric row scaled sqrt(lambda), photometric row scaled
// sqrt(1 - lambda)), inside the same RegistrationICP loop as point-to-plane.  The 3x3 normal equations are solved by an
// unpivoted LDL^T (upstream: Eigen LDLT); a non-positive pivot or fewer than 4 neighbours leaves the gradient at zero.
extern "C" void o3r_pcd_search(const double* pts, int64_t n, const double* q, int64_t nq, int k, double radius, int32_t* idx,
                               double* d2, int32_t* counts);

O3R_API void o3r_color_gradient(const double* xyz, const double* nrm, const double* rgb, int64_t M, double radius, int max_nn,
                                double* grad) {
    std::vector<int32_t> idx((size_t)M * max_nn), cnt((size_t)M);
    o3r_pcd_search(xyz, M, xyz, M, max_nn, radius, idx.data(), nullptr, cnt.data());
    for (int64_t k = 0; k < M; k++) {
        double* g = grad + 3 * k;
        g[0] = g[1] = g[2] = 0.0;
        const int nn = cnt[k];
        if (nn < 4) continue;
        const double* vt = xyz + 3 * k;
        const double* nt = nrm + 3 * k;
        const double it = (rgb[3 * k] + rgb[3 * k + 1] + rgb[3 * k + 2]) / 3.0;
        double AtA[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, Atb[3] = {0, 0, 0};
        for (int i = 1; i <= nn; i++) {
            double row[3], rhs;
            if (i < nn) {
                const int j = idx[(size_t)k * max_nn + i];
                const double* va = xyz + 3 * j;
                const double dn = (va[0] - vt[0]) * nt[0] + (va[1] - vt[1]) * nt[1] + (va[2] - vt[2]) * nt[2];
                for (int c = 0; c < 3; c++) row[c] = (va[c] - dn * nt[c]) - vt[c];
                rhs = (rgb[3 * j] + rgb[3 * j + 1] + rgb[3 * j + 2]) / 3.0 - it;
            } else {
                for (int c = 0; c < 3; c++) row[c] = (double)(nn - 1) * nt[c];
                rhs = 0.0;
            }
            for (int a = 0; a < 3; a++) {
                for (int b = 0; b < 3; b++) AtA[a][b] += row[a] * row[b];
                Atb[a] += row[a] * rhs;
            }
        }
        const double d0 = AtA[0][0];
        if (!(d0 > 0.0)) continue;
        const double l10 = AtA[1][0] / d0, l20 = AtA[2][0] / d0;
        const double d1 = AtA[1][1] - l10 * AtA[1][0];
        if (!(d1 > 0.0)) continue;
        const double l21 = (AtA[2][1] - l20 * AtA[1][0]) / d1;
        const double d2 = AtA[2][2] - l20 * AtA[2][0] - l21 * l21 * d1;
        if (!(d2 > 0.0)) continue;
        const double y0 = Atb[0], y1 = Atb[1] - l10 * y0, y2 = Atb[2] - l20 * y0 - l21 * y1;
        const double z2 = y2 / d2, z1 = y1 / d1 - l21 * z2, z0 = y0 / d0 - l10 * z1 - l20 * z2;
        g[0] = z0; g[1] = z1; g[2] = z2;
    }
}

// Returns the iterations executed, or < 0 on the upstream LogError conditions (-1: max_dist <= 0, -2: no target normals,
// -3: no target colours, -4: no source colours, -5: lambda outside [0, 1]).  grad_out optional [M][3].
O3R_API int o3r_icp_colored(const double* src_xyz, const double* src_rgb, int64_t N, const double* tgt_xyz, const double* tgt_nrm,
                            const double* tgt_rgb, int64_t M, double max_dist, double lambda_geometric, const double* init,
                            int max_iter, double rel_fitness, double rel_rmse, double* out_T, double* out_fitness,
                            double* out_rmse, int32_t* corr_out, double* grad_out) {
    if (max_dist <= 0.0) return -1;
    if (!tgt_nrm) return -2;
    if (!tgt_rgb) return -3;
    if (!src_rgb) return -4;
    if (!(lambda_geometric >= 0.0 && lambda_geometric <= 1.0)) return -5;
    std::vector<double> grad((size_t)3 * M);
    o3r_color_gradient(tgt_xyz, tgt_nrm, tgt_rgb, M, max_dist * 2.0, 30, grad.data());
    if (grad_out) std::memcpy(grad_out, grad.data(), sizeof(double) * 3 * M);
    const double slg = std::sqrt(lambda_geometric), slp = std::sqrt(1.0 - lambda_geometric);
    M4 T;
    std::memcpy(T.m, init, sizeof(T.m));
    std::vector<double> pcd(src_xyz, src_xyz + 3 * N);
    o3r_transform_points(pcd.data(), N, T.m, nullptr);
    Grid g;
    g.build(tgt_xyz, M, max_dist);
    std::vector<int32_t> corr;
    IcpEval res = icp_evaluate(pcd, g, max_dist, corr);
    int it = 0;
    for (; it < max_iter; it++) {
        M4 U = m4_identity();
        if (res.n > 0) {
            NormalEq ne;
            ne.zero();
            for (int64_t i = 0; i < N; i++) {
                const int j = corr[i];
                if (j < 0) continue;
                const double* vs = &pcd[3 * i];
                const double* vt = tgt_xyz + 3 * j;
                const double* nt = tgt_nrm + 3 * j;
                const double* dit = &grad[3 * (size_t)j];
                const double dn = (vs[0] - vt[0]) * nt[0] + (vs[1] - vt[1]) * nt[1] + (vs[2] - vt[2]) * nt[2];
                const double Jg[6] = {slg * (vs[1] * nt[2] - vs[2] * nt[1]), slg * (vs[2] * nt[0] - vs[0] * nt[2]), slg * (vs[0] * nt[1] - vs[1] * nt[0]),
                                      slg * nt[0], slg * nt[1], slg * nt[2]};
                ne.add_row(Jg, slg * dn);
                double proj[3];
                for (int c = 0; c < 3; c++) proj[c] = (vs[c] - dn * nt[c]) - vt[c];
                const double is = (src_rgb[3 * i] + src_rgb[3 * i + 1] + src_rgb[3 * i + 2]) / 3.0;
                const double itg = (tgt_rgb[3 * j] + tgt_rgb[3 * j + 1] + tgt_rgb[3 * j + 2]) / 3.0;
                const double is0 = dit[0] * proj[0] + dit[1] * proj[1] + dit[2] * proj[2] + itg;
                const double dd = dit[0] * nt[0] + dit[1] * nt[1] + dit[2] * nt[2];
                const double m[3] = {slp * (dd * nt[0] - dit[0]), slp * (dd * nt[1] - dit[1]), slp * (dd * nt[2] - dit[2])};   // -dit^T (I - n n^T)
                const double Jp[6] = {vs[1] * m[2] - vs[2] * m[1], vs[2] * m[0] - vs[0] * m[2], vs[0] * m[1] - vs[1] * m[0], m[0], m[1], m[2]};
                ne.add_row(Jp, slp * (is - is0));
            }
            if (!solve_update(ne, &U)) U = m4_identity();
        }
        T = m4_mul(U, T);
        o3r_transform_points(pcd.data(), N, U.m, nullptr);
        IcpEval prev = res;
        res = icp_evaluate(pcd, g, max_dist, corr);
        if (std::fabs(prev.fitness - res.fitness) < rel_fitness && std::fabs(prev.rmse - res.rmse) < rel_rmse) { it++; break; }
    }
    std::memcpy(out_T, T.m, sizeof(T.m));
    *out_fitness = res.fitness;
    *out_rmse = res.rmse;
    if (corr_out) std::memcpy(corr_out, corr.data(), sizeof(int32_t) * N);
    return it;
}

// helpers exposed for tests
O3R_API int o3r_m4_inverse(const double* in, double* out) {
    M4 a, b;
    std::memcpy(a.m, in, sizeof(a.m));
    if (!m4_inverse(a, &b)) return -1;
    std::memcpy(out, b.m, sizeof(b.m));
    return 0;
}
O3R_API int o3r_solve6(const double* A, const double* b, double* x) { return ldlt6_solve(A, b, x) ? 0 : -1; }
O3R_API void o3r_vec6_to_m4(const double* x, double* T) { M4 m = vec6_to_m4(x); std::memcpy(T, m.m, sizeof(m.m)); }
// torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU reference arm of bench.py asks for all host cores explicitly
O3R_API void o3r_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
O3R_API int o3r_num_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
