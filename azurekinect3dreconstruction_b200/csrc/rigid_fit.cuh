// rigid_fit.cuh -- Horn's closed-form absolute orientation (1987): the rigid transform (no scale) minimising
// sum |R s_i + t - t_i|^2, from the largest eigenvector of the 4x4 quaternion matrix (cyclic Jacobi).  Only + - * / sqrt in a
// fixed order, so the CPU restatement (oracle/oracle_pcd.cpp) reproduces it bit for bit.  Upstream: Eigen::umeyama in
// TransformationEstimationPointToPoint::ComputeTransformation -- same optimum, proper rotation by construction.
#pragma once
#include "common.cuh"

// S[a][b] = sum (s_a - ms_a)(t_b - mt_b) (any positive scale), ms / mt = centroids.  T row-major 4x4.
__device__ inline void horn_from_covariance(const double (*S)[3], const double* ms, const double* mt, double* T) {
    double A[4][4], V[4][4];
    A[0][0] = xdadd(xdadd(S[0][0], S[1][1]), S[2][2]);
    A[1][1] = xdsub(xdsub(S[0][0], S[1][1]), S[2][2]);
    A[2][2] = xdsub(xdsub(S[1][1], S[0][0]), S[2][2]);
    A[3][3] = xdsub(xdsub(S[2][2], S[0][0]), S[1][1]);
    A[0][1] = A[1][0] = xdsub(S[1][2], S[2][1]);
    A[0][2] = A[2][0] = xdsub(S[2][0], S[0][2]);
    A[0][3] = A[3][0] = xdsub(S[0][1], S[1][0]);
    A[1][2] = A[2][1] = xdadd(S[0][1], S[1][0]);
    A[1][3] = A[3][1] = xdadd(S[2][0], S[0][2]);
    A[2][3] = A[3][2] = xdadd(S[1][2], S[2][1]);
    for (int a = 0; a < 4; a++) for (int b = 0; b < 4; b++) V[a][b] = a == b ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 16; sweep++) {
        double off = 0.0;
        for (int p = 0; p < 3; p++) for (int q = p + 1; q < 4; q++) off = xdadd(off, xdmul(A[p][q], A[p][q]));
        if (off == 0.0) break;
        for (int p = 0; p < 3; p++)
            for (int q = p + 1; q < 4; q++) {
                const double apq = A[p][q];
                if (apq == 0.0) continue;
                const double theta = xddiv(xdsub(A[q][q], A[p][p]), xdmul(2.0, apq));
                const double tt = xddiv(theta >= 0.0 ? 1.0 : -1.0, xdadd(fabs(theta), sqrt(xdadd(xdmul(theta, theta), 1.0))));
                const double c = xddiv(1.0, sqrt(xdadd(xdmul(tt, tt), 1.0))), sn = xdmul(tt, c);
                for (int k = 0; k < 4; k++) {                 // A <- A J (columns p, q)
                    const double akp = A[k][p], akq = A[k][q];
                    A[k][p] = xdsub(xdmul(c, akp), xdmul(sn, akq));
                    A[k][q] = xdadd(xdmul(sn, akp), xdmul(c, akq));
                }
                for (int k = 0; k < 4; k++) {                 // A <- J^T A (rows p, q)
                    const double apk = A[p][k], aqk = A[q][k];
                    A[p][k] = xdsub(xdmul(c, apk), xdmul(sn, aqk));
                    A[q][k] = xdadd(xdmul(sn, apk), xdmul(c, aqk));
                }
                for (int k = 0; k < 4; k++) {
                    const double vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = xdsub(xdmul(c, vkp), xdmul(sn, vkq));
                    V[k][q] = xdadd(xdmul(sn, vkp), xdmul(c, vkq));
                }
            }
    }
    int best = 0;
    for (int k = 1; k < 4; k++) if (A[k][k] > A[best][best]) best = k;
    double qw = V[0][best], qx = V[1][best], qy = V[2][best], qz = V[3][best];
    const double qn = sqrt(xdadd(xdadd(xdadd(xdmul(qw, qw), xdmul(qx, qx)), xdmul(qy, qy)), xdmul(qz, qz)));
    qw = xddiv(qw, qn); qx = xddiv(qx, qn); qy = xddiv(qy, qn); qz = xddiv(qz, qn);
    double R[3][3];
    R[0][0] = xdsub(1.0, xdmul(2.0, xdadd(xdmul(qy, qy), xdmul(qz, qz))));
    R[0][1] = xdmul(2.0, xdsub(xdmul(qx, qy), xdmul(qw, qz)));
    R[0][2] = xdmul(2.0, xdadd(xdmul(qx, qz), xdmul(qw, qy)));
    R[1][0] = xdmul(2.0, xdadd(xdmul(qx, qy), xdmul(qw, qz)));
    R[1][1] = xdsub(1.0, xdmul(2.0, xdadd(xdmul(qx, qx), xdmul(qz, qz))));
    R[1][2] = xdmul(2.0, xdsub(xdmul(qy, qz), xdmul(qw, qx)));
    R[2][0] = xdmul(2.0, xdsub(xdmul(qx, qz), xdmul(qw, qy)));
    R[2][1] = xdmul(2.0, xdadd(xdmul(qy, qz), xdmul(qw, qx)));
    R[2][2] = xdsub(1.0, xdmul(2.0, xdadd(xdmul(qx, qx), xdmul(qy, qy))));
    for (int a = 0; a < 3; a++) {
        for (int b = 0; b < 3; b++) T[4 * a + b] = R[a][b];
        T[4 * a + 3] = xdsub(mt[a], xdadd(xdadd(xdmul(R[a][0], ms[0]), xdmul(R[a][1], ms[1])), xdmul(R[a][2], ms[2])));
    }
    T[12] = 0.0; T[13] = 0.0; T[14] = 0.0; T[15] = 1.0;
}


__device__ inline void horn_rigid(const double (*s)[3], const double (*t)[3], int n, double* T) {
    double ms[3] = {0, 0, 0}, mt[3] = {0, 0, 0};
    for (int i = 0; i < n; i++)
        for (int c = 0; c < 3; c++) { ms[c] = xdadd(ms[c], s[i][c]); mt[c] = xdadd(mt[c], t[i][c]); }
    for (int c = 0; c < 3; c++) { ms[c] = xddiv(ms[c], (double)n); mt[c] = xddiv(mt[c], (double)n); }
    double S[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < n; i++)
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 3; b++) S[a][b] = xdadd(S[a][b], xdmul(xdsub(s[i][a], ms[a]), xdsub(t[i][b], mt[b])));
    horn_from_covariance(S, ms, mt, T);
}
