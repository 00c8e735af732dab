// K11: LocalGlobalRegistration -- dense point correspondences from the patch score matrices and the local-to-global
// registration on top of them, entirely on the device.
//
// Replaces /root/reference/geotransformer/modules/geotransformer/local_global_registration.py:49-235 (called at
// experiments/*/model.py:198-205 on the output of the optimal transport; use_dustbin = False and
// correspondence_limit = None as in every shipped config, experiments/*/config.py:116-125).  The reference builds
// (B,K,K) helper tensors with scatter writes, calls torch.nonzero, walks the chunk boundaries in a Python loop on the
// host, pads the chunks into a batch with index_put_, and runs weighted Procrustes 6 times, each with a host round trip
// for the SVD.  Here:
//   patch   one CTA per patch pair: exp(scores) in shared memory, row / column top-k as bit masks, the mutual (or
//           either-side) correspondence mask, and that patch pair's own Procrustes hypothesis straight from the masked
//           K x K weights (no padding, no gather)
//   scan    exclusive scan of the per-patch counts
//   emit    correspondences written in (patch, i, j) order -- torch.nonzero's order (:147)
//   inlier  every hypothesis scored on ALL correspondences (:178-183)
//   refine  one CTA: best hypothesis, then num_refinement_steps re-weighted Procrustes solutions (:184-198)
// No host synchronisation; the number of correspondences is left in device memory.
// Equal scores in a top-k are ordered by the smaller index (torch.topk leaves ties unspecified).
#include "geotx_common.cuh"
#include "procrustes.cuh"

namespace geotx {

constexpr int LGR_MAX_K = 128;    // points per patch
constexpr int LGR_MAX_TOPK = 8;
constexpr int LGR_WORDS = LGR_MAX_K / 32;
constexpr float LGR_EPS = 1e-5f;  // WeightedProcrustes default eps (procrustes.py:77)

struct LgrWorkspace {
    unsigned int* corr_bits;  // [P][K][LGR_WORDS]
    int* counts;              // [P]
    int* offsets;             // [P + 1]
    float* transforms;        // [P][16] local hypotheses
    int* inliers;             // [P]  (-1: the patch pair has fewer than correspondence_threshold correspondences)
};

__host__ inline size_t lgr_align(size_t x) { return (x + 255) & ~(size_t)255; }
__host__ inline size_t lgr_layout(int P, int K, char* base, LgrWorkspace* w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += lgr_align(bytes); return p; };
    char* a = take((size_t)P * K * LGR_WORDS * 4);
    char* b = take((size_t)P * 4);
    char* c = take((size_t)(P + 1) * 4);
    char* d = take((size_t)P * 16 * 4);
    char* e = take((size_t)P * 4);
    if (w) {
        w->corr_bits = (unsigned int*)a;
        w->counts = (int*)b;
        w->offsets = (int*)c;
        w->transforms = (float*)d;
        w->inliers = (int*)e;
    }
    return off;
}

// Sums N values per thread over the CTA; every thread gets the totals.  red: [N][32] floats of shared memory.
template <int N>
__device__ __forceinline__ void lgr_block_sum(float (&v)[N], float (*red)[32]) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int k = 0; k < N; ++k) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
    }
    __syncthreads();
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < N; ++k) red[k][wid] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < N; ++k) {
        float s = 0.f;
        for (int w = 0; w < nw; ++w) s += red[k][w];
        v[k] = s;
    }
}

// Weighted Procrustes over the items a CTA can enumerate: `visit(f)` calls f(src, ref, weight) for each item of this
// thread.  procrustes.py:43-59 with weight_thresh = 0.  The transform lands in T (16 floats, shared or global); all
// threads must call it; T is valid after the trailing __syncthreads().
template <class Visit>
__device__ __forceinline__ void lgr_procrustes(Visit visit, float (*red)[32], float* T) {
    float acc[1] = {0.f};
    visit([&](const float* s, const float* r, float w) { acc[0] += w; });
    lgr_block_sum(acc, red);
    const float inv = 1.0f / (acc[0] + LGR_EPS);
    float c[6] = {0, 0, 0, 0, 0, 0};
    visit([&](const float* s, const float* r, float w) {
        const float wn = w * inv;
#pragma unroll
        for (int d = 0; d < 3; ++d) { c[d] += s[d] * wn; c[3 + d] += r[d] * wn; }
    });
    lgr_block_sum(c, red);
    float h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    visit([&](const float* s, const float* r, float w) {
        const float wn = w * inv;
        const float sc[3] = {s[0] - c[0], s[1] - c[1], s[2] - c[2]};
        const float rc[3] = {wn * (r[0] - c[3]), wn * (r[1] - c[4]), wn * (r[2] - c[5])};
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) h[3 * i + j] += sc[i] * rc[j];
    });
    lgr_block_sum(h, red);
    if (threadIdx.x == 0) kabsch_transform(h, c, T);
    __syncthreads();
}

// modules/ops/transformation.py:39 + local_global_registration.py:139-141: |ref - (R src + t)| < radius
__device__ __forceinline__ bool lgr_inlier(const float* T, const float* s, const float* r, float radius) {
    float d2 = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float a = __fmaf_rn(s[2], T[4 * i + 2], __fmaf_rn(s[1], T[4 * i + 1], __fmul_rn(s[0], T[4 * i]))) + T[4 * i + 3];
        const float d = r[i] - a;
        d2 += d * d;
    }
    return sqrtf(d2) < radius;
}

// ---- patch: correspondence mask + local hypothesis of one patch pair -------------------------------------------------
__global__ void __launch_bounds__(256)
lgr_patch_kernel(const float* __restrict__ ref_knn_points, const float* __restrict__ src_knn_points,
                 const uint8_t* __restrict__ ref_masks, const uint8_t* __restrict__ src_masks,
                 const float* __restrict__ score_mat, long long score_bs, long long score_rs,
                 const float* __restrict__ global_scores, int K, int topk, int mutual, float conf_thr, int corr_thr,
                 LgrWorkspace w) {
    extern __shared__ __align__(16) float lgr_S[];  // K x K: exp(score_mat[b])   (:216)
    __shared__ unsigned int rb[LGR_MAX_K][LGR_WORDS], cb[LGR_MAX_K][LGR_WORDS], corr[LGR_MAX_K][LGR_WORDS];
    __shared__ float red[9][32];
    __shared__ int count_s;
    const int b = blockIdx.x, tid = threadIdx.x;
    const float* sm = score_mat + (size_t)b * score_bs;
    for (int e = tid; e < K * K; e += blockDim.x) lgr_S[e] = expf(sm[(size_t)(e / K) * score_rs + (e % K)]);
    for (int e = tid; e < LGR_MAX_K * LGR_WORDS; e += blockDim.x) {
        (&rb[0][0])[e] = 0u;
        (&cb[0][0])[e] = 0u;
    }
    if (tid == 0) count_s = 0;
    __syncthreads();
    // :57-62 / :65-70: the k largest of every row (threads 0..K-1) and of every column (threads K..2K-1), kept if
    // above the confidence threshold.  2K <= 256 threads.
    if (tid < 2 * K) {
        const bool is_row = tid < K;
        const int line = is_row ? tid : tid - K;
        const int stride = is_row ? 1 : K;
        const float* base = lgr_S + (is_row ? line * K : line);
        int taken[LGR_MAX_TOPK];
        for (int t = 0; t < topk; ++t) {
            float best = -__int_as_float(0x7f800000);
            int bi = -1;
            for (int e = 0; e < K; ++e) {
                const float v = base[e * stride];
                bool used = false;
                for (int q = 0; q < t; ++q) used |= taken[q] == e;
                if (!used && v > best) { best = v; bi = e; }
            }
            taken[t] = bi;
            if (bi >= 0 && best > conf_thr) {
                if (is_row) rb[line][bi >> 5] |= 1u << (bi & 31);       // only this thread writes row `line`
                else atomicOr(&cb[bi][line >> 5], 1u << (line & 31));
            }
        }
    }
    __syncthreads();
    // :73-82 merge the two sides, then the validity masks of the patch points
    const uint8_t* rm = ref_masks + (size_t)b * K;
    const uint8_t* sk = src_masks + (size_t)b * K;
    for (int e = tid; e < K * LGR_WORDS; e += blockDim.x) {
        const int i = e / LGR_WORDS, wd = e % LGR_WORDS;
        unsigned int m = mutual ? (rb[i][wd] & cb[i][wd]) : (rb[i][wd] | cb[i][wd]);
        unsigned int cols = 0u;
        for (int j = wd * 32; j < min(K, wd * 32 + 32); ++j) cols |= (sk[j] ? 1u : 0u) << (j & 31);
        m = rm[i] ? (m & cols) : 0u;
        corr[i][wd] = m;
        w.corr_bits[((size_t)b * K + i) * LGR_WORDS + wd] = m;
        if (m) atomicAdd(&count_s, __popc(m));
    }
    __syncthreads();
    const int count = count_s;
    if (tid == 0) {
        w.counts[b] = count;
        w.inliers[b] = count >= corr_thr ? 0 : -1;  // :168-171 chunks below the threshold give no hypothesis
    }
    if (count < corr_thr) return;
    // :173-177 the patch pair's own weighted Procrustes, straight from the masked weights
    const float gs = global_scores ? global_scores[b] : 1.0f;  // :222-223
    const float* rp = ref_knn_points + (size_t)b * K * 3;
    const float* sp = src_knn_points + (size_t)b * K * 3;
    auto visit = [&](auto f) {
        for (int i = tid; i < K; i += blockDim.x)
            for (int wd = 0; wd < LGR_WORDS; ++wd) {
                unsigned int m = corr[i][wd];
                while (m) {
                    const int j = wd * 32 + __ffs(m) - 1;
                    m &= m - 1;
                    f(sp + 3 * j, rp + 3 * i, lgr_S[i * K + j] * gs);
                }
            }
    };
    lgr_procrustes(visit, red, w.transforms + (size_t)b * 16);
}

// ---- scan: where every patch pair's correspondences start ------------------------------------------------------------
__global__ void __launch_bounds__(1024) lgr_scan_kernel(int P, LgrWorkspace w, int* __restrict__ count_out) {
    __shared__ int warp_sum[32];
    __shared__ int carry_s;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < P; base += 1024) {
        const int i = base + threadIdx.x;
        const int c = i < P ? w.counts[i] : 0;
        int incl = c;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += o;
        }
        if (lane == 31) warp_sum[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int s = warp_sum[lane];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, s, off);
                if (lane >= off) s += o;
            }
            warp_sum[lane] = s;
        }
        __syncthreads();
        const int before = carry_s + (wid ? warp_sum[wid - 1] : 0) + incl - c;
        if (i < P) w.offsets[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = before + c;
        __syncthreads();
    }
    if (threadIdx.x == 0) { w.offsets[P] = carry_s; *count_out = carry_s; }
}

// ---- emit: correspondences in (patch, i, j) order (:147-150) -----------------------------------------------------------
__global__ void __launch_bounds__(LGR_MAX_K)
lgr_emit_kernel(const float* __restrict__ ref_knn_points, const float* __restrict__ src_knn_points,
                const float* __restrict__ score_mat, long long score_bs, long long score_rs,
                const float* __restrict__ global_scores, int K, LgrWorkspace w, float* __restrict__ ref_corr,
                float* __restrict__ src_corr, float* __restrict__ corr_scores) {
    __shared__ int row_start[LGR_MAX_K + 1];
    const int b = blockIdx.x, i = threadIdx.x;
    if (w.counts[b] == 0) return;
    unsigned int m[LGR_WORDS];
    int n = 0;
#pragma unroll
    for (int wd = 0; wd < LGR_WORDS; ++wd) {
        m[wd] = i < K ? w.corr_bits[((size_t)b * K + i) * LGR_WORDS + wd] : 0u;
        n += __popc(m[wd]);
    }
    row_start[i + 1] = n;
    if (i == 0) row_start[0] = 0;
    __syncthreads();
    if (i == 0)
        for (int r = 1; r <= LGR_MAX_K; ++r) row_start[r] += row_start[r - 1];
    __syncthreads();
    if (n == 0) return;
    const float gs = global_scores ? global_scores[b] : 1.0f;
    const float* sm = score_mat + (size_t)b * score_bs + (size_t)i * score_rs;
    size_t o = (size_t)w.offsets[b] + row_start[i];
#pragma unroll
    for (int wd = 0; wd < LGR_WORDS; ++wd) {
        unsigned int bits = m[wd];
        while (bits) {
            const int j = wd * 32 + __ffs(bits) - 1;
            bits &= bits - 1;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                ref_corr[3 * o + d] = ref_knn_points[((size_t)b * K + i) * 3 + d];
                src_corr[3 * o + d] = src_knn_points[((size_t)b * K + j) * 3 + d];
            }
            corr_scores[o] = expf(sm[j]) * gs;  // :216, :222-224
            ++o;
        }
    }
}

// ---- inlier: every local hypothesis against all correspondences (:178-183) ----------------------------------------------
__global__ void __launch_bounds__(256)
lgr_inlier_kernel(const float* __restrict__ ref_corr, const float* __restrict__ src_corr, int P, float radius,
                  LgrWorkspace w) {
    __shared__ float T[16];
    __shared__ int total_s;
    const int b = blockIdx.x;
    if (w.inliers[b] < 0) return;
    if (threadIdx.x < 16) T[threadIdx.x] = w.transforms[(size_t)b * 16 + threadIdx.x];
    if (threadIdx.x == 0) total_s = 0;
    __syncthreads();
    const int C = w.offsets[P];
    int n = 0;
    for (int e = threadIdx.x; e < C; e += blockDim.x) n += lgr_inlier(T, src_corr + 3 * (size_t)e, ref_corr + 3 * (size_t)e, radius);
    n = __reduce_add_sync(0xffffffffu, n);
    if ((threadIdx.x & 31) == 0) atomicAdd(&total_s, n);
    __syncthreads();
    if (threadIdx.x == 0) w.inliers[b] = total_s;
}

// ---- refine: best hypothesis, then re-weighted global Procrustes (:184-198) ---------------------------------------------
__global__ void __launch_bounds__(1024)
lgr_refine_kernel(const float* __restrict__ ref_corr, const float* __restrict__ src_corr,
                  const float* __restrict__ corr_scores, int P, float radius, int steps, LgrWorkspace w,
                  float* __restrict__ transform_out) {
    __shared__ float red[9][32];
    __shared__ float T[16];
    __shared__ int best_s;
    const int C = w.offsets[P];
    if (threadIdx.x == 0) {  // :183 argmax over the hypotheses, first maximum
        int best = -1, best_n = -1;
        for (int b = 0; b < P; ++b)
            if (w.inliers[b] > best_n) { best_n = w.inliers[b]; best = b; }
        best_s = best;
    }
    __syncthreads();
    auto all_weighted = [&](auto weight) {
        return [&, weight](auto f) {
            for (int e = threadIdx.x; e < C; e += blockDim.x)
                f(src_corr + 3 * (size_t)e, ref_corr + 3 * (size_t)e, weight(e));
        };
    };
    if (best_s >= 0) {
        if (threadIdx.x < 16) T[threadIdx.x] = w.transforms[(size_t)best_s * 16 + threadIdx.x];
        __syncthreads();
    } else {  // :185-190 no patch pair reached the threshold: start from all correspondences
        lgr_procrustes(all_weighted([&](int e) { return corr_scores[e]; }), red, T);
    }
    for (int s = 0; s < steps; ++s) {  // :184 / :193-198: weights = scores x inliers of the current transform
        float Tc[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) Tc[q] = T[q];
        __syncthreads();  // everybody holds the current transform before thread 0 overwrites T
        lgr_procrustes(all_weighted([&](int e) {
            return lgr_inlier(Tc, src_corr + 3 * (size_t)e, ref_corr + 3 * (size_t)e, radius) ? corr_scores[e] : 0.0f;
        }), red, T);
    }
    if (threadIdx.x < 16) transform_out[threadIdx.x] = T[threadIdx.x];
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
#include "host_util.h"
using namespace geotx;

extern "C" int64_t geotx_lgr_workspace_bytes(int num_patches, int K) {
    if (num_patches <= 0 || K <= 0) return 0;
    return (int64_t)lgr_layout(num_patches, K, nullptr, nullptr);
}

extern "C" int geotx_local_global_registration(
    const float* ref_knn_points, const float* src_knn_points, const uint8_t* ref_knn_masks, const uint8_t* src_knn_masks,
    const float* score_mat, int64_t score_batch_stride, int64_t score_row_stride, const float* global_scores,
    int num_patches, int K, int topk, float acceptance_radius, int mutual, float confidence_threshold,
    int correspondence_threshold, int num_refinement_steps, float* ref_corr_points, float* src_corr_points,
    float* corr_scores, int32_t* count_out, float* transform_out, void* workspace, void* stream) {
    GEOTX_REQUIRE(ref_knn_points && src_knn_points && ref_knn_masks && src_knn_masks && score_mat && ref_corr_points &&
                      src_corr_points && corr_scores && count_out && transform_out && workspace,
                  "geotx_local_global_registration: null pointer");
    GEOTX_REQUIRE(num_patches > 0 && K > 0 && K <= LGR_MAX_K, "geotx_local_global_registration: bad P=%d or K=%d (max %d)",
                  num_patches, K, LGR_MAX_K);
    GEOTX_REQUIRE(topk > 0 && topk <= LGR_MAX_TOPK && topk <= K, "geotx_local_global_registration: bad k=%d (max %d)",
                  topk, LGR_MAX_TOPK);
    GEOTX_REQUIRE(num_refinement_steps >= 1, "geotx_local_global_registration: num_refinement_steps=%d",
                  num_refinement_steps);
    GEOTX_REQUIRE(score_row_stride >= K && score_batch_stride >= (int64_t)K * score_row_stride,
                  "geotx_local_global_registration: score strides %lld/%lld too small for K=%d",
                  (long long)score_batch_stride, (long long)score_row_stride, K);
    GEOTX_REQUIRE((int64_t)num_patches * K * 2 * topk < 2147483647LL,
                  "geotx_local_global_registration: too many correspondences");
    cudaStream_t st = (cudaStream_t)stream;
    LgrWorkspace w;
    lgr_layout(num_patches, K, (char*)workspace, &w);
    const size_t smem = (size_t)K * K * sizeof(float);
    if (smem > 32 * 1024) {
        int rc = check_cuda(cudaFuncSetAttribute(lgr_patch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "geotx_local_global_registration: shared memory");
        if (rc) return rc;
    }
    lgr_patch_kernel<<<num_patches, 256, smem, st>>>(ref_knn_points, src_knn_points, ref_knn_masks, src_knn_masks,
                                                     score_mat, (long long)score_batch_stride,
                                                     (long long)score_row_stride, global_scores, K, topk, mutual,
                                                     confidence_threshold, correspondence_threshold, w);
    lgr_scan_kernel<<<1, 1024, 0, st>>>(num_patches, w, count_out);
    lgr_emit_kernel<<<num_patches, LGR_MAX_K, 0, st>>>(ref_knn_points, src_knn_points, score_mat,
                                                       (long long)score_batch_stride, (long long)score_row_stride,
                                                       global_scores, K, w, ref_corr_points, src_corr_points,
                                                       corr_scores);
    lgr_inlier_kernel<<<num_patches, 256, 0, st>>>(ref_corr_points, src_corr_points, num_patches, acceptance_radius, w);
    lgr_refine_kernel<<<1, 1024, 0, st>>>(ref_corr_points, src_corr_points, corr_scores, num_patches,
                                          acceptance_radius, num_refinement_steps, w, transform_out);
    return launch_status("geotx_local_global_registration");
}
