// K8: batched weighted Procrustes (Kabsch) with the 3x3 SVD on the device (first item).
//
// Replaces /root/reference/geotransformer/modules/registration/procrustes.py:6-73, which is called by the fine
// matching stage (modules/geotransformer/local_global_registration.py:169, :181-191: once per patch correspondence
// batch and then num_refinement_steps times on all correspondences).  The reference copies the batch of 3x3
// covariances to the HOST for torch.svd (procrustes.py:53-54: `H.cpu()`, then `.cuda()` three times) -- the only
// device->host round trips of the inference path, each one a full stream synchronisation.  Here one CTA per problem
// does the two weighted reductions, thread 0 runs a one-sided Jacobi SVD of the 3x3 matrix in fp64 (a few hundred
// flops) and writes the 4x4 transform; nothing leaves the device and nothing synchronises.
//
// Arithmetic follows the reference step by step: weights below weight_thresh are zeroed (:43), normalised by
// (sum + eps) (:44), weighted centroids (:47-48), H = sum_n src_c[n] (w[n] ref_c[n])^T (:52),
// R = V diag(1, 1, sign(det(V U^T))) U^T (:55-57), t = c_ref - R c_src (:59).  R is unique whenever H has rank >= 2
// (sign and ordering conventions of the SVD cancel); for rank-deficient H the reference's result depends on LAPACK's
// choice of null vectors and this kernel completes the bases with cross products (still a proper rotation).
#include "geotx_common.cuh"

namespace geotx {

constexpr int PROC_THREADS = 256;

// Sums K values per thread over the CTA (every thread gets all K totals): one shuffle tree per value, one exchange.
template <int K>
__device__ __forceinline__ void block_sum(float (&v)[K], float (*red)[PROC_THREADS / 32] /*[9][warps]*/) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
    }
    __syncthreads();  // red may still be read from the previous call
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) red[k][threadIdx.x >> 5] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; ++k) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < PROC_THREADS / 32; ++w) s += red[k][w];
        v[k] = s;
    }
}

// One-sided (Hestenes) Jacobi SVD of a 3x3 matrix: A = U diag(sig) V^T, singular values sorted descending,
// columns of U completed to an orthonormal basis when singular values vanish.
__device__ void svd3(const double H[3][3], double U[3][3], double sig[3], double V[3][3]) {
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

__global__ void __launch_bounds__(PROC_THREADS)
weighted_procrustes_kernel(const float* __restrict__ src_points, const float* __restrict__ ref_points,
                           const float* __restrict__ weights, int N, float weight_thresh, float eps,
                           float* __restrict__ transform_out) {
    __shared__ float red[9][PROC_THREADS / 32];
    const int b = blockIdx.x;
    const float* src = src_points + (size_t)b * N * 3;
    const float* ref = ref_points + (size_t)b * N * 3;
    const float* w = weights ? weights + (size_t)b * N : nullptr;
    auto weight = [&](int n) -> float {
        const float x = w ? w[n] : 1.0f;
        return x < weight_thresh ? 0.0f : x;  // procrustes.py:43
    };
    float acc[1] = {0.f};
    for (int n = threadIdx.x; n < N; n += PROC_THREADS) acc[0] += weight(n);
    block_sum(acc, red);
    const float inv = 1.0f / (acc[0] + eps);  // :44
    float c[6] = {0, 0, 0, 0, 0, 0};
    for (int n = threadIdx.x; n < N; n += PROC_THREADS) {
        const float wn = weight(n) * inv;
#pragma unroll
        for (int d = 0; d < 3; ++d) { c[d] += src[3 * n + d] * wn; c[3 + d] += ref[3 * n + d] * wn; }  // :47-48
    }
    block_sum(c, red);
    float h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int n = threadIdx.x; n < N; n += PROC_THREADS) {
        const float wn = weight(n) * inv;
        const float s[3] = {src[3 * n] - c[0], src[3 * n + 1] - c[1], src[3 * n + 2] - c[2]};                 // :49
        const float r[3] = {wn * (ref[3 * n] - c[3]), wn * (ref[3 * n + 1] - c[4]), wn * (ref[3 * n + 2] - c[5])};  // :50,52
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) h[3 * i + j] += s[i] * r[j];
    }
    block_sum(h, red);
    if (threadIdx.x != 0) return;
    double H[3][3], U[3][3], V[3][3], sig[3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) H[i][j] = (double)h[3 * i + j];
    svd3(H, U, sig, V);
    // :55-57  R = V diag(1, 1, sign(det(V U^T))) U^T
    const double dv = det3(V) * det3(U);
    const double sgn = dv > 0 ? 1.0 : (dv < 0 ? -1.0 : 0.0);
    double R[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R[i][j] = V[i][0] * U[j][0] + V[i][1] * U[j][1] + sgn * V[i][2] * U[j][2];
    float* T = transform_out + (size_t)b * 16;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) T[4 * i + j] = (float)R[i][j];
        T[4 * i + 3] = (float)((double)c[3 + i] - (R[i][0] * c[0] + R[i][1] * c[1] + R[i][2] * c[2]));  // :59
    }
    T[12] = 0.f; T[13] = 0.f; T[14] = 0.f; T[15] = 1.f;
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
#include "host_util.h"
using namespace geotx;

extern "C" int geotx_weighted_procrustes(const float* src_points, const float* ref_points, const float* weights, int B,
                                         int N, float weight_thresh, float eps, float* transform_out, void* stream) {
    GEOTX_REQUIRE(src_points && ref_points && transform_out, "geotx_weighted_procrustes: null pointer");
    GEOTX_REQUIRE(B > 0 && N > 0, "geotx_weighted_procrustes: bad sizes B=%d N=%d", B, N);
    weighted_procrustes_kernel<<<B, PROC_THREADS, 0, (cudaStream_t)stream>>>(src_points, ref_points, weights, N,
                                                                            weight_thresh, eps, transform_out);
    return launch_status("geotx_weighted_procrustes");
}
