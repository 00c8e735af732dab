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
//   emit    every CTA sums the counts of the patch pairs before its own and writes its correspondences in
//           (patch, i, j) order -- torch.nonzero's order (:147)
//   inlier  every hypothesis scored on ALL correspondences (:178-183): a thread keeps one correspondence in registers
//           and walks the hypotheses in shared memory
//   refine  one cluster of 8 CTAs: best hypothesis, then num_refinement_steps re-weighted Procrustes solutions
//           (:184-198), the partial moments exchanged through distributed shared memory, one cluster barrier a step
// Every Procrustes solve is ONE pass over its correspondences: the 16 shifted moments (sum w, sum w s', sum w r',
// sum w s' r'^T with s' = s - a, r' = r - b for a fixed pair of points a, b near the data) give the reference's
// normalised weights, centroids and covariance in closed form (lgr_solve); the reference makes three passes.
// No host synchronisation; the number of correspondences is left in device memory.
// Equal scores in a top-k are ordered by the smaller index (torch.topk leaves ties unspecified).
#include "geotx_common.cuh"
#include "procrustes.cuh"
#include "sm100_ptx.cuh"

namespace geotx {

constexpr int LGR_MAX_K = 128;    // points per patch
constexpr int LGR_MAX_TOPK = 8;
constexpr int LGR_WORDS = LGR_MAX_K / 32;
constexpr float LGR_EPS = 1e-5f;  // WeightedProcrustes default eps (procrustes.py:77)

struct LgrWorkspace {
    unsigned int* corr_bits;  // [P][K][LGR_WORDS]
    int* counts;              // [P]
    int* total;               // [1] number of correspondences
    float* transforms;        // [P][16] local hypotheses
    int* inliers;             // [P]  (-1: the patch pair has fewer than correspondence_threshold correspondences)
};

__host__ inline size_t lgr_align(size_t x) { return (x + 255) & ~(size_t)255; }
__host__ inline size_t lgr_layout(int P, int K, char* base, LgrWorkspace* w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += lgr_align(bytes); return p; };
    char* a = take((size_t)P * K * LGR_WORDS * 4);
    char* b = take((size_t)P * 4);
    char* c = take(4);
    char* d = take((size_t)P * 16 * 4);
    char* e = take((size_t)P * 4);
    if (w) {
        w->corr_bits = (unsigned int*)a;
        w->counts = (int*)b;
        w->total = (int*)c;
        w->transforms = (float*)d;
        w->inliers = (int*)e;
    }
    return off;
}


// ---- one-pass weighted Procrustes ----------------------------------------------------------------------------------
// m[0] = sum w, m[1..3] = sum w s', m[4..6] = sum w r', m[7 + 3i + j] = sum w s'_i r'_j   (s' = s - a, r' = r - b)
__device__ __forceinline__ void lgr_accumulate(float (&m)[16], const float* s, const float* r, float w, const float* a,
                                               const float* b) {
    const float sp[3] = {s[0] - a[0], s[1] - a[1], s[2] - a[2]};
    const float rp[3] = {r[0] - b[0], r[1] - b[1], r[2] - b[2]};
    m[0] += w;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float ws = w * sp[i];
        m[1 + i] += ws;
        m[4 + i] = __fmaf_rn(w, rp[i], m[4 + i]);
#pragma unroll
        for (int j = 0; j < 3; ++j) m[7 + 3 * i + j] = __fmaf_rn(ws, rp[j], m[7 + 3 * i + j]);
    }
}

// procrustes.py:43-59 (weight_thresh = 0) from the 16 moments.  With d = W + eps, rho = W / d, S = sum w s' / d,
// R = sum w r' / d, M = sum w s' r'^T / d:  c_src = S + rho a, c_ref = R + rho b, and with u = S - (1 - rho) a,
// v = R - (1 - rho) b:  H = sum w_n (s - c_src)(r - c_ref)^T = M - S v^T - u R^T + rho u v^T.  One thread; T gets the
// 4x4 transform, c (6 floats) the two centroids.
__device__ inline void lgr_solve(const float* m, const float* a, const float* b, float* T, float* c) {
    const double inv = 1.0 / ((double)m[0] + (double)LGR_EPS), rho = (double)m[0] * inv;
    double S[3], R[3], u[3], v[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        S[i] = (double)m[1 + i] * inv;
        R[i] = (double)m[4 + i] * inv;
        u[i] = S[i] - (1.0 - rho) * (double)a[i];
        v[i] = R[i] - (1.0 - rho) * (double)b[i];
    }
    float h[9], cc[6];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            h[3 * i + j] = (float)((double)m[7 + 3 * i + j] * inv - S[i] * v[j] - u[i] * R[j] + rho * u[i] * v[j]);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        cc[i] = (float)(S[i] + rho * (double)a[i]);
        cc[3 + i] = (float)(R[i] + rho * (double)b[i]);
    }
    kabsch_transform<float>(h, cc, T);  // SVD in the reference's precision (torch.svd on an fp32 matrix)
#pragma unroll
    for (int i = 0; i < 6; ++i) c[i] = cc[i];
}

// modules/ops/transformation.py:39 + local_global_registration.py:139-141: |ref - (R src + t)| < radius
// `d2_max` is the largest float whose (correctly rounded) square root is below the radius, found by the launcher, so
// the test is the reference's without the square root.
__device__ __forceinline__ bool lgr_inlier(const float* T, const float* s, const float* r, float d2_max) {
    float d2 = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float a = __fmaf_rn(s[2], T[4 * i + 2], __fmaf_rn(s[1], T[4 * i + 1], __fmul_rn(s[0], T[4 * i]))) + T[4 * i + 3];
        const float d = r[i] - a;
        d2 += d * d;
    }
    return d2 <= d2_max;
}

// Sum of 16 values over a warp in 16 shuffles: every exchange halves the values a lane keeps (offsets 16, 8, 4, 2),
// the last one (offset 1) adds the two lanes that hold the same value.  Returns the total of value
// lgr_warp_value(lane) (both lanes of a pair return it).
__device__ __forceinline__ int lgr_warp_value(int lane) {
    return ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
}
__device__ __forceinline__ float lgr_warp_sum(float (&m)[16], int lane) {
#pragma unroll
    for (int half = 8; half >= 1; half >>= 1) {
        const bool upper = (lane & (half << 1)) != 0;  // this lane keeps m[half .. 2 half), its partner m[0 .. half)
#pragma unroll
        for (int k = 0; k < half; ++k) {
            const float send = upper ? m[k] : m[k + half], keep = upper ? m[k + half] : m[k];
            m[k] = keep + __shfl_xor_sync(0xffffffffu, send, half << 1);
        }
    }
    return m[0] + __shfl_xor_sync(0xffffffffu, m[0], 1);
}

// ---- patch: correspondence mask + local hypothesis of one patch pair -------------------------------------------------
constexpr int LGR_PATCH_THREADS = 256;
__host__ __device__ inline int lgr_pitch(int K) { return K | 1; }  // odd row pitch: rows AND columns conflict-free
__host__ inline size_t lgr_patch_smem(int K) { return ((size_t)K * lgr_pitch(K) + 6 * (size_t)K) * sizeof(float); }

__global__ void __launch_bounds__(LGR_PATCH_THREADS)
lgr_patch_kernel(const float* __restrict__ ref_knn_points, const float* __restrict__ src_knn_points,
                 const uint8_t* __restrict__ ref_masks, const uint8_t* __restrict__ src_masks,
                 const float* __restrict__ score_mat, long long score_bs, long long score_rs,
                 const float* __restrict__ global_scores, int K, int topk, int mutual, float conf_thr, int corr_thr,
                 LgrWorkspace w) {
    extern __shared__ __align__(16) float lgr_smem[];
    const int pitch = lgr_pitch(K);
    float* S = lgr_smem;           // K x pitch: exp(score_mat[b])   (:216)
    float* pr = S + K * pitch;     // K x 3 ref patch points
    float* ps = pr + 3 * K;        // K x 3 src patch points
    __shared__ unsigned int rb[LGR_MAX_K][LGR_WORDS], cb[LGR_MAX_K][LGR_WORDS], corr[LGR_MAX_K][LGR_WORDS];
    __shared__ float red[16][LGR_PATCH_THREADS / 32], tot_s[16];
    __shared__ int count_s, first_s;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float* sm = score_mat + (size_t)b * score_bs;
    // all loads of a batch in flight before the first exp: the matrix comes from HBM once, latency paid K*K/2048 times
    const int total = K * K;
    for (int base = tid; base < total; base += LGR_PATCH_THREADS * 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int e = base + u * LGR_PATCH_THREADS;
            v[u] = e < total ? sm[(size_t)(e / K) * score_rs + (e % K)] : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int e = base + u * LGR_PATCH_THREADS;
            if (e < total) S[(e / K) * pitch + (e % K)] = expf(v[u]);
        }
    }
    for (int e = tid; e < 3 * K; e += LGR_PATCH_THREADS) {
        pr[e] = ref_knn_points[(size_t)b * K * 3 + e];
        ps[e] = src_knn_points[(size_t)b * K * 3 + e];
    }
    for (int e = tid; e < LGR_MAX_K * LGR_WORDS; e += LGR_PATCH_THREADS) {
        (&rb[0][0])[e] = 0u;
        (&cb[0][0])[e] = 0u;
    }
    if (tid == 0) { count_s = 0; first_s = 0x7fffffff; }
    __syncthreads();
    // :57-62 / :65-70: the k largest of every row (threads 0..K-1) and of every column (threads K..2K-1), kept if
    // above the confidence threshold.  2K <= 256 threads.  Pass t takes the successor of pass t-1 in (value descending,
    // index ascending) order, so nothing has to be remembered but the last pick.
    if (tid < 2 * K) {
        const bool is_row = tid < K;
        const int line = is_row ? tid : tid - K;
        const int stride = is_row ? 1 : pitch;
        const float* base = S + (is_row ? line * pitch : line);
        float pv = __int_as_float(0x7f800000);
        int pe = -1;
        for (int t = 0; t < topk; ++t) {
            float best = -__int_as_float(0x7f800000);
            int bi = -1;
#pragma unroll 4
            for (int e = 0; e < K; ++e) {
                const float v = base[e * stride];
                const bool cand = v < pv || (v == pv && e > pe);
                if (cand && v > best) { best = v; bi = e; }
            }
            if (bi < 0 || !(best > conf_thr)) break;  // the rest of the order is not larger
            pv = best;
            pe = bi;
            if (is_row) rb[line][bi >> 5] |= 1u << (bi & 31);  // only this thread writes row `line`
            else atomicOr(&cb[bi][line >> 5], 1u << (line & 31));
        }
    }
    __syncthreads();
    // :73-82 merge the two sides, then the validity masks of the patch points
    const uint8_t* rm = ref_masks + (size_t)b * K;
    const uint8_t* sk = src_masks + (size_t)b * K;
    for (int e = tid; e < K * LGR_WORDS; e += LGR_PATCH_THREADS) {
        const int i = e / LGR_WORDS, wd = e % LGR_WORDS;
        unsigned int m = mutual ? (rb[i][wd] & cb[i][wd]) : (rb[i][wd] | cb[i][wd]);
        if (m && rm[i]) {
            unsigned int cols = 0u;
            for (unsigned int rest = m; rest;) {
                const int j = wd * 32 + __ffs(rest) - 1;
                rest &= rest - 1;
                cols |= (sk[j] ? 1u : 0u) << (j & 31);
            }
            m &= cols;
        } else {
            m = 0u;
        }
        corr[i][wd] = m;
        w.corr_bits[((size_t)b * K + i) * LGR_WORDS + wd] = m;
        if (m) {
            atomicAdd(&count_s, __popc(m));
            atomicMin(&first_s, i * LGR_MAX_K + wd * 32 + __ffs(m) - 1);
        }
    }
    __syncthreads();
    const int count = count_s;
    if (tid == 0) {
        w.counts[b] = count;
        w.inliers[b] = (count >= corr_thr && count > 0) ? 0 : -1;  // :168-171 small chunks give no hypothesis
    }
    if (count < corr_thr || count == 0) return;
    // :173-177 the patch pair's own weighted Procrustes, straight from the masked weights, shifted to the first
    // correspondence's points
    const float gs = global_scores ? global_scores[b] : 1.0f;  // :222-223
    const float* a = ps + 3 * (first_s % LGR_MAX_K);
    const float* bb = pr + 3 * (first_s / LGR_MAX_K);
    float m16[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) m16[q] = 0.f;
    for (int e = tid; e < K * LGR_WORDS; e += LGR_PATCH_THREADS) {
        const int i = e / LGR_WORDS, wd = e % LGR_WORDS;
        for (unsigned int rest = corr[i][wd]; rest;) {
            const int j = wd * 32 + __ffs(rest) - 1;
            rest &= rest - 1;
            lgr_accumulate(m16, ps + 3 * j, pr + 3 * i, S[i * pitch + j] * gs, a, bb);
        }
    }
    {
        const float t = lgr_warp_sum(m16, lane);
        if ((lane & 1) == 0) red[lgr_warp_value(lane)][wid] = t;
    }
    __syncthreads();
    if (tid < 16) {
        float t = 0.f;
#pragma unroll
        for (int ww = 0; ww < LGR_PATCH_THREADS / 32; ++ww) t += red[tid][ww];
        tot_s[tid] = t;
    }
    __syncthreads();
    if (tid == 0) {
        float c[6];
        lgr_solve(tot_s, a, bb, w.transforms + (size_t)b * 16, c);
    }
}

// ---- emit: correspondences in (patch, i, j) order (:147-150) -----------------------------------------------------------
__global__ void __launch_bounds__(LGR_MAX_K)
lgr_emit_kernel(const float* __restrict__ ref_knn_points, const float* __restrict__ src_knn_points,
                const float* __restrict__ score_mat, long long score_bs, long long score_rs,
                const float* __restrict__ global_scores, int P, int K, LgrWorkspace w, float* __restrict__ ref_corr,
                float* __restrict__ src_corr, float* __restrict__ corr_scores, int* __restrict__ count_out) {
    __shared__ int row_start[LGR_MAX_K + 1];
    __shared__ int warp_part[LGR_MAX_K / 32];
    const int b = blockIdx.x, i = threadIdx.x, lane = i & 31, wid = i >> 5;
    // where this patch pair's correspondences start: the counts of the pairs before it
    int before = 0;
    for (int p = i; p < b; p += LGR_MAX_K) before += w.counts[p];
    before = __reduce_add_sync(0xffffffffu, before);
    if (lane == 0) warp_part[wid] = before;
    unsigned int m[LGR_WORDS];
    int n = 0;
#pragma unroll
    for (int wd = 0; wd < LGR_WORDS; ++wd) {
        m[wd] = i < K ? w.corr_bits[((size_t)b * K + i) * LGR_WORDS + wd] : 0u;
        n += __popc(m[wd]);
    }
    // exclusive scan of the row counts over the CTA
    int incl = n;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += o;
    }
    if (lane == 31) row_start[wid] = incl;
    __syncthreads();
    int offset = 0, row0 = incl - n;
#pragma unroll
    for (int ww = 0; ww < LGR_MAX_K / 32; ++ww) {
        offset += warp_part[ww];
        if (ww < wid) row0 += row_start[ww];
    }
    if (b == P - 1 && i == LGR_MAX_K - 1) {  // the last thread of the last pair knows the total
        *w.total = offset + row0 + n;
        *count_out = offset + row0 + n;
    }
    if (n == 0) return;
    const float gs = global_scores ? global_scores[b] : 1.0f;
    const float* sm = score_mat + (size_t)b * score_bs + (size_t)i * score_rs;
    size_t o = (size_t)offset + row0;
    const float r0 = ref_knn_points[((size_t)b * K + i) * 3], r1 = ref_knn_points[((size_t)b * K + i) * 3 + 1],
                r2 = ref_knn_points[((size_t)b * K + i) * 3 + 2];
#pragma unroll
    for (int wd = 0; wd < LGR_WORDS; ++wd) {
        unsigned int bits = m[wd];
        while (bits) {
            const int j = wd * 32 + __ffs(bits) - 1;
            bits &= bits - 1;
            const float* sp = src_knn_points + ((size_t)b * K + j) * 3;
            ref_corr[3 * o] = r0; ref_corr[3 * o + 1] = r1; ref_corr[3 * o + 2] = r2;
            src_corr[3 * o] = sp[0]; src_corr[3 * o + 1] = sp[1]; src_corr[3 * o + 2] = sp[2];
            corr_scores[o] = expf(sm[j]) * gs;  // :216, :222-224
            ++o;
        }
    }
}

// ---- inlier: every local hypothesis against all correspondences (:178-183) ----------------------------------------------
constexpr int LGR_INL_THREADS = 256;  // correspondences per tile, one per thread
constexpr int LGR_INL_HYPS = 16;      // hypotheses per CTA (grid.y walks them): enough CTAs to fill the SMs

__global__ void __launch_bounds__(LGR_INL_THREADS)
lgr_inlier_kernel(const float* __restrict__ ref_corr, const float* __restrict__ src_corr, int P, float d2_max,
                  LgrWorkspace w) {
    __shared__ __align__(16) float Ts[LGR_INL_HYPS][12];
    __shared__ int valid[LGR_INL_HYPS], hits[LGR_INL_HYPS];
    const int C = *w.total;
    const int h0 = blockIdx.y * LGR_INL_HYPS, nh = min(LGR_INL_HYPS, P - h0);
    if ((long long)blockIdx.x * LGR_INL_THREADS >= C) return;
    const int tid = threadIdx.x, lane = tid & 31;
    for (int e = tid; e < nh; e += LGR_INL_THREADS) {
        valid[e] = w.inliers[h0 + e] >= 0;  // stable: only hypotheses that start at 0 are ever added to
        hits[e] = 0;
    }
    for (int e = tid; e < nh * 12; e += LGR_INL_THREADS)
        Ts[e / 12][e % 12] = w.transforms[(size_t)(h0 + e / 12) * 16 + e % 12];
    __syncthreads();
    for (long long tile = blockIdx.x; tile * LGR_INL_THREADS < C; tile += gridDim.x) {
        const long long e = tile * LGR_INL_THREADS + tid;
        const bool live = e < C;
        float s[3] = {0.f, 0.f, 0.f}, r[3] = {0.f, 0.f, 0.f};
        if (live) {
#pragma unroll
            for (int d = 0; d < 3; ++d) { s[d] = src_corr[3 * e + d]; r[d] = ref_corr[3 * e + d]; }
        }
#pragma unroll 4
        for (int h = 0; h < nh; ++h) {
            if (!valid[h]) continue;
            const unsigned int ballot = __ballot_sync(0xffffffffu, live && lgr_inlier(Ts[h], s, r, d2_max));
            if (lane == 0 && ballot) atomicAdd(&hits[h], __popc(ballot));
        }
    }
    __syncthreads();
    for (int e = tid; e < nh; e += LGR_INL_THREADS)
        if (hits[e]) atomicAdd(&w.inliers[h0 + e], hits[e]);
}

// ---- refine: best hypothesis, then re-weighted global Procrustes (:184-198) ---------------------------------------------
constexpr int LGR_REF_CTAS = 8;  // one portable cluster
constexpr int LGR_REF_THREADS = 512;
constexpr int LGR_REF_SLOTS = 4;  // correspondences a thread keeps in registers: 16384 over the cluster

__global__ void __cluster_dims__(LGR_REF_CTAS, 1, 1) __launch_bounds__(LGR_REF_THREADS)
lgr_refine_kernel(const float* __restrict__ ref_corr, const float* __restrict__ src_corr,
                  const float* __restrict__ corr_scores, int P, float d2_max, int steps, LgrWorkspace w,
                  float* __restrict__ transform_out) {
    __shared__ float red[16][LGR_REF_THREADS / 32];
    __shared__ __align__(16) float part[2][16];  // this CTA's partial moments, by step parity (peers read them)
    __shared__ float tot[16], T[16], shift[6];
    __shared__ unsigned long long best_part[LGR_REF_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const unsigned int rank = ptx::cluster_ctarank();
    const int C = *w.total;
    // :183 argmax of the inlier counts over the hypotheses, first maximum (every CTA finds it for itself)
    unsigned long long key = 0ull;
    for (int p = tid; p < P; p += LGR_REF_THREADS) {
        const int n = w.inliers[p];
        if (n >= 0) key = max(key, ((unsigned long long)(unsigned int)(n + 1) << 32) | (0xffffffffu - (unsigned int)p));
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) key = max(key, __shfl_xor_sync(0xffffffffu, key, off));
    if (lane == 0) best_part[wid] = key;
    __syncthreads();
#pragma unroll
    for (int ww = 0; ww < LGR_REF_THREADS / 32; ++ww) key = max(key, best_part[ww]);
    const int best = key ? (int)(0xffffffffu - (unsigned int)(key & 0xffffffffull)) : -1;
    if (tid < 16) T[tid] = best >= 0 ? w.transforms[(size_t)best * 16 + tid] : (tid % 5 == 0 ? 1.f : 0.f);
    if (tid < 3) {  // the moments are taken about the first correspondence, later about the last centroids
        shift[tid] = C > 0 ? src_corr[tid] : 0.f;
        shift[3 + tid] = C > 0 ? ref_corr[tid] : 0.f;
    }
    // :185-190 without a hypothesis the first solve takes all correspondences with their scores; every solve after
    // it (and every solve of the normal path, :184 / :193-198) weights them by scores x inliers of the last transform
    const int iters = steps + (best < 0 ? 1 : 0);
    // the correspondences of this thread stay in registers over the steps (absent ones have weight 0)
    constexpr int stride = LGR_REF_CTAS * LGR_REF_THREADS;
    const int e0 = rank * LGR_REF_THREADS + tid;
    float cs[LGR_REF_SLOTS][3], cr[LGR_REF_SLOTS][3], cw[LGR_REF_SLOTS];
#pragma unroll
    for (int q = 0; q < LGR_REF_SLOTS; ++q) {
        const int e = e0 + q * stride;
        const bool live = e < C;
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            cs[q][d] = live ? src_corr[3 * (size_t)e + d] : 0.f;
            cr[q][d] = live ? ref_corr[3 * (size_t)e + d] : 0.f;
        }
        cw[q] = live ? corr_scores[e] : 0.f;
    }
    for (int it = 0; it < iters; ++it) {
        __syncthreads();  // T and shift of the last solve are in place
        const bool plain = best < 0 && it == 0;
        float Tl[16], a[3], b[3], m[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) { Tl[q] = T[q]; m[q] = 0.f; }
#pragma unroll
        for (int d = 0; d < 3; ++d) { a[d] = shift[d]; b[d] = shift[3 + d]; }
#pragma unroll
        for (int q = 0; q < LGR_REF_SLOTS; ++q)
            lgr_accumulate(m, cs[q], cr[q], (plain || lgr_inlier(Tl, cs[q], cr[q], d2_max)) ? cw[q] : 0.f, a, b);
        for (int e = e0 + LGR_REF_SLOTS * stride; e < C; e += stride) {  // what the register slots do not hold
            const float s[3] = {src_corr[3 * (size_t)e], src_corr[3 * (size_t)e + 1], src_corr[3 * (size_t)e + 2]};
            const float r[3] = {ref_corr[3 * (size_t)e], ref_corr[3 * (size_t)e + 1], ref_corr[3 * (size_t)e + 2]};
            const float sc = corr_scores[e];
            lgr_accumulate(m, s, r, (plain || lgr_inlier(Tl, s, r, d2_max)) ? sc : 0.f, a, b);
        }
        {
            const float t = lgr_warp_sum(m, lane);
            if ((lane & 1) == 0) red[lgr_warp_value(lane)][wid] = t;
        }
        __syncthreads();
        if (tid < 16) {
            float t = 0.f;
#pragma unroll
            for (int ww = 0; ww < LGR_REF_THREADS / 32; ++ww) t += red[tid][ww];
            part[it & 1][tid] = t;
        }
        ptx::cluster_sync();  // every CTA's partial of this step is visible cluster-wide
        if (tid < 16) {
            const uint32_t mine = (uint32_t)__cvta_generic_to_shared(&part[it & 1][tid]);
            float t = 0.f;
#pragma unroll
            for (unsigned int z = 0; z < LGR_REF_CTAS; ++z) {  // fixed order: all CTAs get the same bits
                float x;
                asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(x) : "r"(ptx::map_to_cta(mine, z)));
                t += x;
            }
            tot[tid] = t;
        }
        __syncthreads();
        if (tid == 0) lgr_solve(tot, shift, shift + 3, T, shift);  // reads a, b before it writes the new centroids
    }
    ptx::cluster_sync();  // nobody leaves while a peer may still read its partials
    if (rank == 0 && tid < 16) transform_out[tid] = T[tid];
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
#include <math.h>
#include "host_util.h"
using namespace geotx;

// sqrt(d2) < radius  <=>  d2 <= lgr_d2_max(radius): sqrtf is correctly rounded and monotone on both sides.
static float lgr_d2_max(float radius) {
    if (!(radius > 0.f)) return -1.f;
    float x = radius * radius;
    while (sqrtf(x) >= radius) x = nextafterf(x, 0.f);
    while (sqrtf(nextafterf(x, INFINITY)) < radius) x = nextafterf(x, INFINITY);
    return x;
}

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
    const int64_t cap = (int64_t)num_patches * K * 2 * topk;
    GEOTX_REQUIRE(cap < 2147483647LL, "geotx_local_global_registration: too many correspondences");
    cudaStream_t st = (cudaStream_t)stream;
    LgrWorkspace w;
    lgr_layout(num_patches, K, (char*)workspace, &w);
    const size_t smem = lgr_patch_smem(K);
    if (smem > 32 * 1024) {
        int rc = check_cuda(cudaFuncSetAttribute(lgr_patch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "geotx_local_global_registration: shared memory");
        if (rc) return rc;
    }
    lgr_patch_kernel<<<num_patches, LGR_PATCH_THREADS, smem, st>>>(
        ref_knn_points, src_knn_points, ref_knn_masks, src_knn_masks, score_mat, (long long)score_batch_stride,
        (long long)score_row_stride, global_scores, K, topk, mutual, confidence_threshold, correspondence_threshold, w);
    lgr_emit_kernel<<<num_patches, LGR_MAX_K, 0, st>>>(ref_knn_points, src_knn_points, score_mat,
                                                       (long long)score_batch_stride, (long long)score_row_stride,
                                                       global_scores, num_patches, K, w, ref_corr_points,
                                                       src_corr_points, corr_scores, count_out);
    const int tiles = (int)((cap + LGR_INL_THREADS - 1) / LGR_INL_THREADS);
    const dim3 inl_grid(tiles < sm_count() ? tiles : sm_count(), (num_patches + LGR_INL_HYPS - 1) / LGR_INL_HYPS);
    const float d2_max = lgr_d2_max(acceptance_radius);
    lgr_inlier_kernel<<<inl_grid, LGR_INL_THREADS, 0, st>>>(ref_corr_points, src_corr_points, num_patches, d2_max, w);
    lgr_refine_kernel<<<LGR_REF_CTAS, LGR_REF_THREADS, 0, st>>>(ref_corr_points, src_corr_points, corr_scores,
                                                                num_patches, d2_max, num_refinement_steps,
                                                                w, transform_out);
    return launch_status("geotx_local_global_registration");
}
