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
#include "procrustes.cuh"

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
    if (threadIdx.x == 0) kabsch_transform(h, c, transform_out + (size_t)b * 16);  // :53-59
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
