// 3x3 SVD and the Kabsch solution from accumulated moments: shared by K8 (weighted Procrustes) and K11
// (local-to-global registration).  modules/registration/procrustes.py:52-59.
#pragma once
#include <cuda_runtime.h>

namespace geotx {

// Tolerances of the Jacobi iteration and of the rank decision, per scalar type.
template <typename F> struct SvdTol;
template <> struct SvdTol<double> {
    static __device__ __forceinline__ double skip2() { return 1e-34; }  // rotation skipped: gamma^2 <= skip2 alpha beta
    static __device__ __forceinline__ double done2() { return 1e-30; }  // sweep converged:  gamma^2 <  done2 alpha beta
    static __device__ __forceinline__ double rank() { return 1e-12; }
    static __device__ __forceinline__ double floor() { return 1e-300; }
};
template <> struct SvdTol<float> {
    static __device__ __forceinline__ float skip2() { return 1e-16f; }
    static __device__ __forceinline__ float done2() { return 4e-14f; }
    static __device__ __forceinline__ float rank() { return 1e-6f; }
    static __device__ __forceinline__ float floor() { return 1e-37f; }
};

// One-sided (Hestenes) Jacobi SVD of a 3x3 matrix: A = U diag(sig) V^T, singular values sorted descending,
// columns of U completed to an orthonormal basis when singular values vanish.  All loops over matrix indices unroll,
// so the matrices live in registers; a rotation costs two square roots and two divisions.
template <typename F>
__device__ inline void svd3(const F H[3][3], F U[3][3], F sig[3], F V[3][3]) {
    F A[3][3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) { A[r][c] = H[r][c]; V[r][c] = r == c ? F(1) : F(0); }
    for (int sweep = 0; sweep < 30; ++sweep) {
        bool done = true;
#pragma unroll
        for (int p = 0; p < 2; ++p)
#pragma unroll
            for (int q = p + 1; q < 3; ++q) {
                F alpha = 0, beta = 0, gamma = 0;
#pragma unroll
                for (int r = 0; r < 3; ++r) { alpha += A[r][p] * A[r][p]; beta += A[r][q] * A[r][q]; gamma += A[r][p] * A[r][q]; }
                const F g2 = gamma * gamma, ab = alpha * beta;
                if (g2 <= SvdTol<F>::skip2() * ab) continue;
                if (!(g2 < SvdTol<F>::done2() * ab)) done = false;
                // tan of the rotation angle: t = sign(zeta) / (|zeta| + sqrt(1 + zeta^2)), zeta = (beta - alpha) / (2 gamma)
                const F d = beta - alpha, g = gamma + gamma;
                const F t = (d >= 0 ? g : -g) / (fabs(d) + sqrt(d * d + g * g));
                const F c = F(1) / sqrt(F(1) + t * t), s = c * t;
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    const F ap = A[r][p], aq = A[r][q];
                    A[r][p] = c * ap - s * aq;
                    A[r][q] = s * ap + c * aq;
                    const F vp = V[r][p], vq = V[r][q];
                    V[r][p] = c * vp - s * vq;
                    V[r][q] = s * vp + c * vq;
                }
            }
        if (done) break;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) sig[c] = sqrt(A[0][c] * A[0][c] + A[1][c] * A[1][c] + A[2][c] * A[2][c]);
    // sort descending (columns of A and V together)
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2 - a; ++b)
            if (sig[b] < sig[b + 1]) {
                F tmp = sig[b]; sig[b] = sig[b + 1]; sig[b + 1] = tmp;
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    tmp = A[r][b]; A[r][b] = A[r][b + 1]; A[r][b + 1] = tmp;
                    tmp = V[r][b]; V[r][b] = V[r][b + 1]; V[r][b + 1] = tmp;
                }
            }
    const F tiny = SvdTol<F>::rank() * fmax(sig[0], SvdTol<F>::floor());
    int rank = 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        if (sig[c] > tiny) {
            const F inv = F(1) / sig[c];
#pragma unroll
            for (int r = 0; r < 3; ++r) U[r][c] = A[r][c] * inv;
            ++rank;
        }
    }
    if (rank == 0) {
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c) U[r][c] = r == c ? F(1) : F(0);
    } else {
        if (rank == 1) {  // any unit vector orthogonal to u0: e_k - (e_k . u0) u0 for the smallest |u0[k]|
            const F a0 = fabs(U[0][0]), a1 = fabs(U[1][0]), a2 = fabs(U[2][0]);
            const int k = a0 <= a1 ? (a0 <= a2 ? 0 : 2) : (a1 <= a2 ? 1 : 2);
            const F d = k == 0 ? U[0][0] : (k == 1 ? U[1][0] : U[2][0]);
            F n2 = 0;
#pragma unroll
            for (int r = 0; r < 3; ++r) { U[r][1] = (r == k ? F(1) : F(0)) - d * U[r][0]; n2 += U[r][1] * U[r][1]; }
            n2 = F(1) / sqrt(n2);
#pragma unroll
            for (int r = 0; r < 3; ++r) U[r][1] *= n2;
        }
        if (rank <= 2) {  // u2 = u0 x u1
            U[0][2] = U[1][0] * U[2][1] - U[2][0] * U[1][1];
            U[1][2] = U[2][0] * U[0][1] - U[0][0] * U[2][1];
            U[2][2] = U[0][0] * U[1][1] - U[1][0] * U[0][1];
        }
    }
}

template <typename F>
__device__ __forceinline__ F det3(const F M[3][3]) {
    return M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0]) +
           M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
}

// procrustes.py:53-59 from the 3x3 covariance h (row-major, src x ref) and the two weighted centroids c = (src, ref):
// R = V diag(1, 1, sign(det(V U^T))) U^T, t = c_ref - R c_src; T is the row-major 4x4 [R t; 0 0 0 1].  One thread.
// F is the scalar type of the SVD: double in K8, float (the reference's own precision, torch.svd on fp32) in K11.
template <typename F = double>
__device__ inline void kabsch_transform(const float* h, const float* c, float* T) {
    F H[3][3], U[3][3], V[3][3], sig[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) H[i][j] = (F)h[3 * i + j];
    svd3<F>(H, U, sig, V);
    const F dv = det3<F>(V) * det3<F>(U);
    const F sgn = dv > 0 ? F(1) : (dv < 0 ? F(-1) : F(0));
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        F R[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            R[j] = V[i][0] * U[j][0] + V[i][1] * U[j][1] + sgn * V[i][2] * U[j][2];
            T[4 * i + j] = (float)R[j];
        }
        T[4 * i + 3] = (float)((F)c[3 + i] - (R[0] * (F)c[0] + R[1] * (F)c[1] + R[2] * (F)c[2]));
    }
    T[12] = 0.f; T[13] = 0.f; T[14] = 0.f; T[15] = 1.f;
}

}  // namespace geotx
