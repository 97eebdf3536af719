// host_math.h -- tiny f64 host-side linear algebra for the C-ABI layer (row-major 4x4).
// The 4x4 inverse follows the cofactor closed form (what Eigen uses for fixed-size 4x4, SURVEY A.3:
// camera_pose = extrinsic.inverse()).  Compiled with -ffp-contract=off so the operation sequence
// rounds identically wherever it is built.
#pragma once
#include <cmath>
#include <cstring>

namespace b2r {

inline bool m4_inverse(const double* m, double* out) {
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
    for (int i = 0; i < 16; i++) out[i] = inv[i] * idet;
    return true;
}

inline void m4_mul(const double* a, const double* b, double* r) {
    double t[16];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            double s = 0.0;
            for (int k = 0; k < 4; k++) s += a[i * 4 + k] * b[k * 4 + j];
            t[i * 4 + j] = s;
        }
    std::memcpy(r, t, sizeof(t));
}

inline void m4_identity(double* r) {
    for (int i = 0; i < 16; i++) r[i] = (i % 5 == 0) ? 1.0 : 0.0;
}

// a6 colour-to-depth gather constants (the chain is defined in unproject.cu / oracle.cpp "a6"): folded in f64,
// rounded once to f32.  dK, cK = {fx, fy, cx, cy}; d2c = row-major rigid transform (>= 12 doubles).
// out16 = {ax, bx, ay, by, M[3][4]} with M = [[Fx,0,Cx+0.5],[0,Fy,Cy+0.5],[0,0,1]] * [R|t].
inline void gather_constants(const double* dK, const double* cK, const double* d2c, float* out16) {
    out16[0] = (float)(1.0 / dK[0]); out16[1] = (float)(-dK[2] / dK[0]);
    out16[2] = (float)(1.0 / dK[1]); out16[3] = (float)(-dK[3] / dK[1]);
    const double cu = cK[2] + 0.5, cv = cK[3] + 0.5;
    for (int j = 0; j < 4; j++) {
        out16[4 + j] = (float)(cK[0] * d2c[j] + cu * d2c[8 + j]);
        out16[8 + j] = (float)(cK[1] * d2c[4 + j] + cv * d2c[8 + j]);
        out16[12 + j] = (float)d2c[8 + j];
    }
}

}  // namespace b2r
