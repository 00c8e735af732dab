// K12: patch matching scores -- the glue between the coarse matching (K5) and the optimal transport (K7).
//
// Replaces experiments/*/model.py:169-170,176-179,187-188: for every node correspondence p the features of the K points
// of its ref patch and of its src patch are gathered from the (zero-padded) fine feature matrices and multiplied,
//     scores[p, i, j] = <feats_ref[ref_knn[ref_corr[p], i]], feats_src[src_knn[src_corr[p], j]]> / sqrt(C).
// The reference concatenates a zero row to both feature matrices (two full copies), gathers two (P, K, C) tensors
// (2 x 16 MB at the 3DMatch size), runs a batched einsum and a division: seven launches and ~100 MB of traffic for
// 0.5 GFLOP.  Here one CTA per correspondence gathers its 2K rows straight into shared memory (channel-major, so that
// the inner product reads are conflict-free 128-bit loads), 64 channels at a time, and keeps the K x K tile in
// registers; nothing but the indices and the result touches global memory.  fp32 FMA throughout (the reference's
// einsum is fp32; no tensor cores: a TF32 product would move the transport plan by ~1e-2).
#include "geotx_common.cuh"

namespace geotx {

constexpr int PS_THREADS = 256;
constexpr int PS_CHUNK = 64;  // channels staged per step

template <int KP>  // patch size rounded up to 64
__global__ void __launch_bounds__(PS_THREADS)
patch_scores_kernel(const float* __restrict__ ref_feats, long long n_ref, const float* __restrict__ src_feats,
                    long long n_src, int C, const long long* __restrict__ ref_knn, const long long* __restrict__ src_knn,
                    int K, const long long* __restrict__ ref_corr, const long long* __restrict__ src_corr,
                    float divisor, float* __restrict__ scores) {
    constexpr int PITCH = KP + 4;  // floats per channel row: 16-byte aligned, the two halves of a warp on disjoint banks
    constexpr int NB = KP / 64;    // 64 x 64 output blocks per side
    extern __shared__ __align__(16) float ps_smem[];
    float* At = ps_smem;                    // [PS_CHUNK][PITCH]  ref patch, channel-major
    float* Bt = At + PS_CHUNK * PITCH;      // [PS_CHUNK][PITCH]  src patch
    __shared__ long long rows_a[KP], rows_b[KP];
    const int p = blockIdx.x, tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    {
        const long long na = ref_corr[p], nb = src_corr[p];
        for (int r = tid; r < KP; r += PS_THREADS) {
            rows_a[r] = r < K ? ref_knn[na * K + r] : n_ref;  // n_ref / n_src: the zero row of the padded matrices
            rows_b[r] = r < K ? src_knn[nb * K + r] : n_src;
        }
    }
    float acc[NB][NB][4][4];
#pragma unroll
    for (int a = 0; a < NB; ++a)
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int s = 0; s < 4; ++s) acc[a][b][r][s] = 0.f;
    for (int c0 = 0; c0 < C; c0 += PS_CHUNK) {
        __syncthreads();  // rows_* written / the previous chunk consumed
        // a lane pair reads the two halves of a 32-byte sector of one row; 16 rows per warp instruction
        for (int u = tid; u < KP * (PS_CHUNK / 4); u += PS_THREADS) {
            const int half = u & 1, r = (u >> 1) % KP, c = ((u >> 1) / KP) * 8 + half * 4;
            const long long ga = rows_a[r], gb = rows_b[r];
            float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
            if (c0 + c + 3 < C) {
                if (ga < n_ref) va = *reinterpret_cast<const float4*>(ref_feats + ga * C + c0 + c);
                if (gb < n_src) vb = *reinterpret_cast<const float4*>(src_feats + gb * C + c0 + c);
            }
            At[(c + 0) * PITCH + r] = va.x; At[(c + 1) * PITCH + r] = va.y;
            At[(c + 2) * PITCH + r] = va.z; At[(c + 3) * PITCH + r] = va.w;
            Bt[(c + 0) * PITCH + r] = vb.x; Bt[(c + 1) * PITCH + r] = vb.y;
            Bt[(c + 2) * PITCH + r] = vb.z; Bt[(c + 3) * PITCH + r] = vb.w;
        }
        __syncthreads();
#pragma unroll 4
        for (int c = 0; c < PS_CHUNK; ++c) {
            float4 a4[NB], b4[NB];
#pragma unroll
            for (int a = 0; a < NB; ++a) a4[a] = *reinterpret_cast<const float4*>(At + c * PITCH + a * 64 + ty * 4);
#pragma unroll
            for (int b = 0; b < NB; ++b) b4[b] = *reinterpret_cast<const float4*>(Bt + c * PITCH + b * 64 + tx * 4);
#pragma unroll
            for (int a = 0; a < NB; ++a)
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    const float av[4] = {a4[a].x, a4[a].y, a4[a].z, a4[a].w};
                    const float bv[4] = {b4[b].x, b4[b].y, b4[b].z, b4[b].w};
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int s = 0; s < 4; ++s) acc[a][b][r][s] = __fmaf_rn(av[r], bv[s], acc[a][b][r][s]);
                }
        }
    }
    float* out = scores + (size_t)p * K * K;
#pragma unroll
    for (int a = 0; a < NB; ++a)
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int i = a * 64 + ty * 4 + r;
#pragma unroll
                for (int s = 0; s < 4; ++s) {
                    const int j = b * 64 + tx * 4 + s;
                    if (i < K && j < K) out[(size_t)i * K + j] = __fdiv_rn(acc[a][b][r][s], divisor);  // model.py:188
                }
            }
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
#include <math.h>
#include "host_util.h"
using namespace geotx;

template <int KP>
static int launch_patch_scores(const float* ref_feats, int64_t n_ref, const float* src_feats, int64_t n_src, int C,
                               const int64_t* ref_knn, const int64_t* src_knn, int K, const int64_t* ref_corr,
                               const int64_t* src_corr, int P, float* scores, cudaStream_t st) {
    const size_t smem = 2 * (size_t)PS_CHUNK * (KP + 4) * sizeof(float);
    if (smem > 48 * 1024) {
        int rc = check_cuda(cudaFuncSetAttribute(patch_scores_kernel<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)smem), "geotx_patch_scores: shared memory");
        if (rc) return rc;
    }
    const float divisor = (float)sqrt((double)C);  // model.py:188: feats_f.shape[1] ** 0.5, a Python float
    patch_scores_kernel<KP><<<P, PS_THREADS, smem, st>>>(ref_feats, (long long)n_ref, src_feats, (long long)n_src, C,
                                                         (const long long*)ref_knn, (const long long*)src_knn, K,
                                                         (const long long*)ref_corr, (const long long*)src_corr,
                                                         divisor, scores);
    return launch_status("geotx_patch_scores");
}

extern "C" int geotx_patch_scores(const float* ref_feats, int64_t n_ref, const float* src_feats, int64_t n_src, int C,
                                  const int64_t* ref_node_knn_indices, const int64_t* src_node_knn_indices, int K,
                                  const int64_t* ref_node_corr_indices, const int64_t* src_node_corr_indices, int P,
                                  float* scores, void* stream) {
    GEOTX_REQUIRE(ref_feats && src_feats && ref_node_knn_indices && src_node_knn_indices && ref_node_corr_indices &&
                      src_node_corr_indices && scores, "geotx_patch_scores: null pointer");
    GEOTX_REQUIRE(P > 0 && K > 0 && K <= 128 && C > 0 && C % 4 == 0 && n_ref >= 0 && n_src >= 0,
                  "geotx_patch_scores: bad sizes P=%d K=%d (max 128) C=%d (multiple of 4)", P, K, C);
    cudaStream_t st = (cudaStream_t)stream;
    if (K <= 64)
        return launch_patch_scores<64>(ref_feats, n_ref, src_feats, n_src, C, ref_node_knn_indices, src_node_knn_indices,
                                       K, ref_node_corr_indices, src_node_corr_indices, P, scores, st);
    return launch_patch_scores<128>(ref_feats, n_ref, src_feats, n_src, C, ref_node_knn_indices, src_node_knn_indices, K,
                                    ref_node_corr_indices, src_node_corr_indices, P, scores, st);
}
