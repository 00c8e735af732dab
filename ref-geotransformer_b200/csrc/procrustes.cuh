// 3x3 SVD and the Kabsch solution from accumulated moments: shared by K8 (weighted Procrustes) and K11
// (local-to-global registration).  modules/registration/procrustes.py:52-59.
#pragma once
#include <cuda_runtime.h>

namespace geotx {

// One-sided (Hestenes) Jacobi SVD of a 3x3 matrix: A = U diag(sig) V^T, singular values sorted descending,
// columns of U completed to an orthonormal basis when singular values vanish.
__device__ inline void svd3(const double H[3][3], double U[3][3], double sig[3], double V[3][3]) {
    double A[3][3];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) { A[r][c] = H[r][c]; V[r][c] = r == c ? 1.0 : 0.0; }
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double alpha = 0, beta = 0, gamma = 0;
                for (int r = 0; r < 3; ++r) { alpha += A[r][p] * A[r][p]; beta += A[r][q] * A[r][q]; gamma += A[r][p] * A[r][q]; }
                if (gamma == 0.0 || fabs(gamma) <= 1e-17 * sqrt(alpha * beta)) continue;
                off = fmax(off, fabs(gamma) / sqrt(alpha * beta));
                const double zeta = (beta - alpha) / (2.0 * gamma);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int r = 0; r < 3; ++r) {
                    const double ap = A[r][p], aq = A[r][q];
                    A[r][p] = c * ap - s * aq;
                    A[r][q] = s * ap + c * aq;
                    const double vp = V[r][p], vq = V[r][q];
                    V[r][p] = c * vp - s * vq;
                    V[r][q] = s * vp + c * vq;
                }
            }
        if (off < 1e-15) break;
    }
    for (int c = 0; c < 3; ++c) sig[c] = sqrt(A[0][c] * A[0][c] + A[1][c] * A[1][c] + A[2][c] * A[2][c]);
    // sort descending (columns of A and V together)
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2 - a; ++b)
            if (sig[b] < sig[b + 1]) {
                double tmp = sig[b]; sig[b] = sig[b + 1]; sig[b + 1] = tmp;
                for (int r = 0; r < 3; ++r) {
                    tmp = A[r][b]; A[r][b] = A[r][b + 1]; A[r][b + 1] = tmp;
                    tmp = V[r][b]; V[r][b] = V[r][b + 1]; V[r][b + 1] = tmp;
                }
            }
    const double tiny = 1e-12 * fmax(sig[0], 1e-300);
    int rank = 0;
    for (int c = 0; c < 3; ++c) {
        if (sig[c] > tiny) { for (int r = 0; r < 3; ++r) U[r][c] = A[r][c] / sig[c]; ++rank; }
    }
    if (rank == 0) {
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) U[r][c] = r == c ? 1.0 : 0.0;
    } else {
        if (rank == 1) {  // any unit vector orthogonal to u0
            int k = fabs(U[0][0]) <= fabs(U[1][0]) ? (fabs(U[0][0]) <= fabs(U[2][0]) ? 0 : 2)
                                                   : (fabs(U[1][0]) <= fabs(U[2][0]) ? 1 : 2);
            double e[3] = {0, 0, 0};
            e[k] = 1.0;
            const double d = U[k][0];
            double n2 = 0;
            for (int r = 0; r < 3; ++r) { U[r][1] = e[r] - d * U[r][0]; n2 += U[r][1] * U[r][1]; }
            n2 = sqrt(n2);
            for (int r = 0; r < 3; ++r) U[r][1] /= n2;
        }
        if (rank <= 2) {  // u2 = u0 x u1
            U[0][2] = U[1][0] * U[2][1] - U[2][0] * U[1][1];
            U[1][2] = U[2][0] * U[0][1] - U[0][0] * U[2][1];
            U[2][2] = U[0][0] * U[1][1] - U[1][0] * U[0][1];
        }
    }
}

__device__ __forceinline__ double det3(const double M[3][3]) {
    return M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0]) +
           M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
}

// procrustes.py:53-59 from the 3x3 covariance h (row-major, src x ref) and the two weighted centroids c = (src, ref):
// R = V diag(1, 1, sign(det(V U^T))) U^T, t = c_ref - R c_src; T is the row-major 4x4 [R t; 0 0 0 1].  One thread.
__device__ inline void kabsch_transform(const float* h, const float* c, float* T) {
    double H[3][3], U[3][3], V[3][3], sig[3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) H[i][j] = (double)h[3 * i + j];
    svd3(H, U, sig, V);
    const double dv = det3(V) * det3(U);
    const double sgn = dv > 0 ? 1.0 : (dv < 0 ? -1.0 : 0.0);
    double R[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R[i][j] = V[i][0] * U[j][0] + V[i][1] * U[j][1] + sgn * V[i][2] * U[j][2];
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) T[4 * i + j] = (float)R[i][j];
        T[4 * i + 3] = (float)((double)c[3 + i] - (R[i][0] * c[0] + R[i][1] * c[1] + R[i][2] * c[2]));
    }
    T[12] = 0.f; T[13] = 0.f; T[14] = 0.f; T[15] = 1.f;
}

}  // namespace geotx
