// K5: coarse matching behind the hot path.
//
// Replaces, for one registration pair,
//   /root/reference/experiments/*/model.py:141-142            F.normalize(feats_c, p=2, dim=1) on both clouds
//   /root/reference/geotransformer/modules/geotransformer/superpoint_matching.py:13-51
//       exp(-pairwise_distance(normalized=True)), dual normalisation by row and column sums, GLOBAL top-K
//   /root/reference/geotransformer/modules/ops/pairwise_distance.py:24-25,30 (2 - 2 xy, clamp at 0)
// The reference materialises five (n,m) fp32 matrices (xy, exp, two quotients, product) and sorts all n*m entries
// in torch.topk.  Here one (n,m) fp32 buffer is written once (scores, then normalised in place, L2-resident at the
// BASELINE sizes) and the top-K is an exact radix SELECT on 64-bit keys (score bits, inverted flat index), so ties
// are broken by the lowest flat index and the result does not depend on scheduling:
//   spm_prepare  : unit rows (optional), workspace reset
//   spm_scores   : fp32 FFMA tiles of xy -> exp(-max(0, 2 - 2xy)), masked entries 0; per-tile row / column partial
//                  sums written to fixed slots (no float atomics: sums are reproducible)
//   spm_sums     : partials added in tile order; K_eff = min(K, valid_rows * valid_cols)
//   spm_pass<p>  : p = 0 also rewrites the buffer as (s / rowsum) * (s / colsum) (IEEE divisions, the reference's
//                  operation order); every pass histograms one 10/11-bit digit of the keys that match the digits
//                  already fixed; the last CTA to finish (ticket) fixes the next digit.  Passes after the one that
//                  isolates exactly K_eff keys return immediately.
//   spm_collect  : appends every key >= threshold (exactly K_eff of them); the last CTA sorts them (bitonic, shared
//                  memory) and writes (ref index, src index, score) in descending order.
// All arithmetic is fp32 on the CUDA cores on purpose: the job is 2 n m D flop (0.5 GFLOP at n = m = 1024, D = 256),
// and the top-K ORDER is decided at the 1e-6 relative level, out of reach of tf32 / half tensor-core inputs.
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {

constexpr int SPM_TILE = 64;        // scores tile edge
constexpr int SPM_KC = 16;          // contraction chunk
constexpr int SPM_PITCH = 68;       // floats per staged k-row (16-byte aligned, 2-way conflicts on the transposing stores only)
constexpr int SPM_BINS = 2048;      // histogram bins (widest digit: 11 bits)
constexpr int SPM_PASSES = 6;       // digits of the 64-bit key, most significant first
constexpr int SPM_MAX_K = 4096;     // keys the final shared-memory sort holds
constexpr uint32_t SPM_MASKED = 0xFFFFFFFFu;  // bit pattern stored for a masked (removed) entry; never a real score

__device__ __constant__ int c_spm_shift[SPM_PASSES] = {53, 42, 32, 21, 10, 0};
__device__ __constant__ int c_spm_width[SPM_PASSES] = {11, 11, 10, 11, 11, 10};

struct SpmState {
    unsigned long long prefix;  // digits fixed so far (lower bits zero); finally the K_eff-th largest key or a lower bound of its bucket
    unsigned int need;          // keys still to take from the bucket `prefix` names
    unsigned int resolved;      // 1: keys >= prefix are exactly the K_eff selected ones
    unsigned int k_eff;
    unsigned int list_count;
    unsigned int ticket[SPM_PASSES + 1];
    unsigned int pad[3];
};

struct SpmWorkspace {  // carved out of the caller's scratch, see spm_layout()
    SpmState* state;
    unsigned int* hist;  // [SPM_PASSES][SPM_BINS]
    float* row_sum;      // [n]
    float* col_sum;      // [m]
    float* row_part;     // [tiles_m][n]
    float* col_part;     // [tiles_n][m]
    float* ref_unit;     // [n][d]
    float* src_unit;     // [m][d]
    float* scores;       // [n][m]
    unsigned long long* list;  // [num_corr]
    size_t bytes;
};

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static SpmWorkspace spm_layout(void* base, int n, int m, int d, int num_corr) {
    SpmWorkspace w;
    const int tiles_n = (n + SPM_TILE - 1) / SPM_TILE, tiles_m = (m + SPM_TILE - 1) / SPM_TILE;
    char* p = (char*)base;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char* r = p + off;
        off += align256(bytes);
        return (void*)r;
    };
    w.state = (SpmState*)take(sizeof(SpmState));
    w.hist = (unsigned int*)take(sizeof(unsigned int) * SPM_PASSES * SPM_BINS);
    w.row_sum = (float*)take(sizeof(float) * n);
    w.col_sum = (float*)take(sizeof(float) * m);
    w.row_part = (float*)take(sizeof(float) * (size_t)tiles_m * n);
    w.col_part = (float*)take(sizeof(float) * (size_t)tiles_n * m);
    w.ref_unit = (float*)take(sizeof(float) * (size_t)n * d);
    w.src_unit = (float*)take(sizeof(float) * (size_t)m * d);
    w.scores = (float*)take(sizeof(float) * (size_t)n * m);
    w.list = (unsigned long long*)take(sizeof(unsigned long long) * (size_t)num_corr);
    w.bytes = off;
    return w;
}

// ---------------------------------------------------------------------------------------------------------------
// spm_prepare: one warp per feature row.  F.normalize(p=2, dim=1, eps=1e-12): x / max(||x||_2, eps).
// Block 0 also clears the selection state and the histograms of this call.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
spm_prepare_kernel(const float* __restrict__ ref, const float* __restrict__ src, int n, int m, int d, int normalize,
                   float* __restrict__ ref_unit, float* __restrict__ src_unit, SpmState* state,
                   unsigned int* hist) {
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < SPM_PASSES * SPM_BINS; i += blockDim.x) hist[i] = 0u;
        if (threadIdx.x < sizeof(SpmState) / 4) ((unsigned int*)state)[threadIdx.x] = 0u;
    }
    if (!normalize) return;
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= n + m) return;
    const float* x = row < n ? ref + (size_t)row * d : src + (size_t)(row - n) * d;
    float* y = row < n ? ref_unit + (size_t)row * d : src_unit + (size_t)(row - n) * d;
    float ss = 0.f;
    for (int c = lane; c < d; c += 32) {
        const float v = __ldg(x + c);
        ss = fmaf(v, v, ss);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
    const float denom = fmaxf(sqrtf(ss), 1e-12f);
    for (int c = lane; c < d; c += 32) y[c] = __fdiv_rn(__ldg(x + c), denom);
}

// ---------------------------------------------------------------------------------------------------------------
// spm_scores: 64 x 64 tile per CTA, 256 threads, 4 x 4 outputs per thread, k chunks of 16 staged transposed in shared
// memory with a register prefetch of the next chunk.  Dot products accumulate in ascending k with FMA.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
spm_scores_kernel(const float* __restrict__ a, const float* __restrict__ b, int n, int m, int d,
                  const uint8_t* __restrict__ ref_keep, const uint8_t* __restrict__ src_keep, int dual,
                  float* __restrict__ scores, float* __restrict__ row_part, float* __restrict__ col_part) {
    __shared__ __align__(16) float As[2][SPM_KC][SPM_PITCH];
    __shared__ __align__(16) float Bs[2][SPM_KC][SPM_PITCH];
    __shared__ float s_col[16][SPM_TILE];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int i0 = blockIdx.y * SPM_TILE, j0 = blockIdx.x * SPM_TILE;
    const int lrow = tid >> 2, lk = (tid & 3) * 4;  // this thread stages row lrow, k offsets lk..lk+3 of each operand
    const bool a_ok = i0 + lrow < n, b_ok = j0 + lrow < m;
    const float* ap = a + (size_t)(a_ok ? i0 + lrow : 0) * d;
    const float* bp = b + (size_t)(b_ok ? j0 + lrow : 0) * d;
    const bool vec = (d & 3) == 0;
    auto fetch = [&](const float* p, bool ok, int k) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok) {
            if (vec && k + 3 < d) {
                v = __ldg((const float4*)(p + k));
            } else {
                if (k < d) v.x = __ldg(p + k);
                if (k + 1 < d) v.y = __ldg(p + k + 1);
                if (k + 2 < d) v.z = __ldg(p + k + 2);
                if (k + 3 < d) v.w = __ldg(p + k + 3);
            }
        }
        return v;
    };
    auto stage = [&](float (*S)[SPM_PITCH], const float4& v) {
        S[lk][lrow] = v.x;
        S[lk + 1][lrow] = v.y;
        S[lk + 2][lrow] = v.z;
        S[lk + 3][lrow] = v.w;
    };
    float acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;
    const int chunks = (d + SPM_KC - 1) / SPM_KC;
    float4 ra = fetch(ap, a_ok, lk), rb = fetch(bp, b_ok, lk);
    stage(As[0], ra);
    stage(Bs[0], rb);
    __syncthreads();
    for (int ch = 0; ch < chunks; ++ch) {
        const int cur = ch & 1;
        if (ch + 1 < chunks) {
            ra = fetch(ap, a_ok, (ch + 1) * SPM_KC + lk);
            rb = fetch(bp, b_ok, (ch + 1) * SPM_KC + lk);
        }
#pragma unroll
        for (int kk = 0; kk < SPM_KC; ++kk) {
            const float4 av = *(const float4*)&As[cur][kk][ty * 4];
            const float4 bv = *(const float4*)&Bs[cur][kk][tx * 4];
            const float ar[4] = {av.x, av.y, av.z, av.w}, br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(ar[r], br[c], acc[r][c]);
        }
        if (ch + 1 < chunks) {
            stage(As[cur ^ 1], ra);
            stage(Bs[cur ^ 1], rb);
        }
        __syncthreads();
    }
    // exp(-clamp(2 - 2 xy, 0)); removed rows / columns contribute 0 to the sums (superpoint_matching.py:32-37)
    bool keep_r[4], keep_c[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int i = i0 + ty * 4 + r;
        keep_r[r] = i < n && (!ref_keep || ref_keep[i]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const int j = j0 + tx * 4 + c;
        keep_c[c] = j < m && (!src_keep || src_keep[j]);
    }
    float rs[4] = {0.f, 0.f, 0.f, 0.f}, cs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float d2 = fmaxf(2.0f - 2.0f * acc[r][c], 0.0f);
            const float s = (keep_r[r] && keep_c[c]) ? expf(-d2) : 0.f;
            acc[r][c] = s;
            rs[r] += s;
            cs[c] += s;
        }
        const int i = i0 + ty * 4 + r;
        if (i < n) {
            float* out = scores + (size_t)i * m + j0 + tx * 4;
            if ((m & 3) == 0 && j0 + tx * 4 + 3 < m) {
                *(float4*)out = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
            } else {
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (j0 + tx * 4 + c < m) out[c] = acc[r][c];
            }
        }
    }
    if (!dual) return;
    // row partials: the 16 threads of a row group are 16 consecutive lanes -> fixed xor tree
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) rs[r] += __shfl_xor_sync(0xffffffffu, rs[r], off);
        const int i = i0 + ty * 4 + r;
        if (tx == 0 && i < n) row_part[(size_t)blockIdx.x * n + i] = rs[r];
    }
    // column partials: through shared memory, added in ascending ty
#pragma unroll
    for (int c = 0; c < 4; ++c) s_col[ty][tx * 4 + c] = cs[c];
    __syncthreads();
    if (tid < SPM_TILE && j0 + tid < m) {
        float t = 0.f;
#pragma unroll
        for (int y = 0; y < 16; ++y) t += s_col[y][tid];
        col_part[(size_t)blockIdx.y * m + j0 + tid] = t;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// spm_sums: row / column sums from the tile partials (ascending tile order); block 0 also counts the kept rows and
// columns and fixes K_eff = min(num_corr, kept_rows * kept_cols) (superpoint_matching.py:42).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
spm_sums_kernel(const float* __restrict__ row_part, const float* __restrict__ col_part, int n, int m, int tiles_n,
                int tiles_m, int dual, const uint8_t* __restrict__ ref_keep, const uint8_t* __restrict__ src_keep,
                int num_corr, float* __restrict__ row_sum, float* __restrict__ col_sum, SpmState* state) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (dual) {
        if (t < n) {
            float s = 0.f;
            for (int k = 0; k < tiles_m; ++k) s += row_part[(size_t)k * n + t];
            row_sum[t] = s;
        } else if (t < n + m) {
            const int j = t - n;
            float s = 0.f;
            for (int k = 0; k < tiles_n; ++k) s += col_part[(size_t)k * m + j];
            col_sum[j] = s;
        }
    }
    if (blockIdx.x == 0) {
        __shared__ unsigned int s_cnt[2];
        if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0u;
        __syncthreads();
        unsigned int r = 0, c = 0;
        for (int i = threadIdx.x; i < n; i += blockDim.x) r += (!ref_keep || ref_keep[i]) ? 1u : 0u;
        for (int j = threadIdx.x; j < m; j += blockDim.x) c += (!src_keep || src_keep[j]) ? 1u : 0u;
        atomicAdd(&s_cnt[0], r);
        atomicAdd(&s_cnt[1], c);
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned long long total = (unsigned long long)s_cnt[0] * s_cnt[1];
            const unsigned int k_eff = (unsigned int)(total < (unsigned long long)num_corr ? total : (unsigned long long)num_corr);
            state->k_eff = k_eff;
            state->need = k_eff;
            state->prefix = 0ull;
            if (k_eff == 0) {  // nothing to select: an unreachable threshold
                state->prefix = ~0ull;
                state->resolved = 1u;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Radix select.  key = (score bits << 32) | (0xFFFFFFFF - flat index): unique, larger = better, ties on the score go
// to the lower flat index.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long spm_key(uint32_t bits, uint32_t idx) {
    return ((unsigned long long)bits << 32) | (unsigned long long)(0xFFFFFFFFu - idx);
}

// Last CTA of a pass: walk the digit histogram from the top until `need` keys are covered.
__device__ void spm_fix_digit(SpmState* state, const unsigned int* hist, int pass, unsigned int* s_part) {
    const int tid = threadIdx.x;  // 256 threads, 8 bins each, bin 2047 first
    unsigned int local = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) local += __ldcg(hist + (SPM_BINS - 1 - (tid * 8 + b)));
    s_part[tid] = local;
    __syncthreads();
    if (tid == 0) {
        const unsigned int need = state->need;
        unsigned int cum = 0;
        int g = 0;
        while (g < 255 && cum + s_part[g] < need) cum += s_part[g++];
        int bin = SPM_BINS - 1 - g * 8;
        unsigned int h = __ldcg(hist + bin);
        while (bin > 0 && cum + h < need) {
            cum += h;
            --bin;
            h = __ldcg(hist + bin);
        }
        state->prefix |= (unsigned long long)bin << c_spm_shift[pass];
        if (cum + h <= need) {  // the bucket is taken whole (== need; < need cannot happen while K_eff <= candidates)
            state->resolved = 1u;
        } else {
            state->need = need - cum;
        }
    }
}

template <bool FIRST>
__global__ void __launch_bounds__(256)
spm_pass_kernel(float* __restrict__ scores, int n, int m, int dual, const float* __restrict__ row_sum,
                const float* __restrict__ col_sum, const uint8_t* __restrict__ ref_keep,
                const uint8_t* __restrict__ src_keep, int pass, SpmState* state, unsigned int* hist_all) {
    __shared__ unsigned int s_hist[SPM_BINS];
    __shared__ unsigned int s_last;
    if (!FIRST && state->resolved) return;  // written by an earlier kernel of the stream: uniform
    unsigned int* hist = hist_all + pass * SPM_BINS;
    for (int i = threadIdx.x; i < SPM_BINS; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();
    const int shift = c_spm_shift[pass], width = c_spm_width[pass];
    const unsigned long long prefix = FIRST ? 0ull : state->prefix;
    const int above = shift + width;  // bits above this digit must equal the prefix
    const unsigned int dmask = (1u << width) - 1u;
    const size_t total = (size_t)n * m;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        uint32_t bits;
        if (FIRST) {
            const int i = (int)(idx / (unsigned)m), j = (int)(idx - (size_t)i * m);
            const bool keep = (!ref_keep || ref_keep[i]) && (!src_keep || src_keep[j]);
            float f = scores[idx];
            if (dual) f = __fdiv_rn(f, __ldg(row_sum + i)) * __fdiv_rn(f, __ldg(col_sum + j));  // superpoint_matching.py:39-41
            bits = keep ? __float_as_uint(f) : SPM_MASKED;
            ((uint32_t*)scores)[idx] = bits;
        } else {
            bits = ((const uint32_t*)scores)[idx];
        }
        if (bits == SPM_MASKED) continue;
        const unsigned long long key = spm_key(bits, (uint32_t)idx);
        if (FIRST || (above < 64 && (key >> above) == (prefix >> above)))
            atomicAdd(&s_hist[(unsigned int)(key >> shift) & dmask], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < SPM_BINS; i += blockDim.x) {
        const unsigned int v = s_hist[i];
        if (v) atomicAdd(hist + i, v);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&state->ticket[pass], 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (s_last) {
        __threadfence();
        spm_fix_digit(state, hist, pass, s_hist);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// spm_collect: keys >= threshold are the K_eff winners; the last CTA sorts them and writes the three outputs.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
spm_collect_kernel(const float* __restrict__ scores, int n, int m, SpmState* state, unsigned long long* list,
                   int num_corr, int64_t* __restrict__ ref_idx, int64_t* __restrict__ src_idx,
                   float* __restrict__ score_out, int32_t* __restrict__ count_out) {
    extern __shared__ unsigned long long s_keys[];
    __shared__ unsigned int s_last;
    const unsigned long long threshold = state->prefix;
    const unsigned int k_eff = state->k_eff;
    const size_t total = (size_t)n * m;
    if (k_eff > 0) {
        for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
            const uint32_t bits = ((const uint32_t*)scores)[idx];
            if (bits == SPM_MASKED) continue;
            const unsigned long long key = spm_key(bits, (uint32_t)idx);
            if (key >= threshold) {
                const unsigned int pos = atomicAdd(&state->list_count, 1u);
                if (pos < (unsigned int)num_corr) list[pos] = key;
            }
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&state->ticket[SPM_PASSES], 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    int size = 1;
    while (size < (int)k_eff) size <<= 1;
    for (int i = threadIdx.x; i < size; i += blockDim.x) s_keys[i] = i < (int)k_eff ? __ldcg(list + i) : 0ull;
    __syncthreads();
    for (int k = 2; k <= size; k <<= 1) {          // bitonic sort, descending
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < size; i += blockDim.x) {
                const int l = i ^ j;
                if (l > i) {
                    const unsigned long long a = s_keys[i], b = s_keys[l];
                    const bool desc = (i & k) == 0;
                    if (desc ? a < b : a > b) {
                        s_keys[i] = b;
                        s_keys[l] = a;
                    }
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < num_corr; i += blockDim.x) {
        if (i < (int)k_eff) {
            const unsigned long long key = s_keys[i];
            const uint32_t idx = 0xFFFFFFFFu - (uint32_t)key;
            const uint32_t r = idx / (uint32_t)m;
            ref_idx[i] = (int64_t)r;
            src_idx[i] = (int64_t)(idx - r * (uint32_t)m);
            score_out[i] = __uint_as_float((uint32_t)(key >> 32));
        } else {
            ref_idx[i] = -1;
            src_idx[i] = -1;
            score_out[i] = 0.f;
        }
    }
    if (threadIdx.x == 0) *count_out = (int32_t)k_eff;
}

}  // namespace geotx

using namespace geotx;

extern "C" int64_t geotx_spm_workspace_bytes(int n, int m, int d, int num_corr) {
    if (n <= 0 || m <= 0 || d <= 0 || num_corr <= 0) return 0;
    return (int64_t)spm_layout(nullptr, n, m, d, num_corr).bytes;
}

extern "C" int geotx_superpoint_matching(const float* ref_feats, const float* src_feats, int n, int m, int d,
                                         const uint8_t* ref_keep, const uint8_t* src_keep, int normalize,
                                         float* ref_unit_out, float* src_unit_out, int dual_normalization,
                                         int num_corr, int64_t* ref_idx_out, int64_t* src_idx_out, float* score_out,
                                         int32_t* count_out, float* score_matrix, void* workspace, void* stream) {
    GEOTX_REQUIRE(ref_feats && src_feats && ref_idx_out && src_idx_out && score_out && count_out && workspace,
                  "geotx_superpoint_matching: null pointer");
    GEOTX_REQUIRE(n > 0 && m > 0 && d > 0, "geotx_superpoint_matching: empty input (n=%d m=%d d=%d)", n, m, d);
    GEOTX_REQUIRE((unsigned long long)n * (unsigned long long)m < 0xFFFFFFFFull,
                  "geotx_superpoint_matching: n*m = %llu does not fit the 32-bit flat index",
                  (unsigned long long)n * (unsigned long long)m);
    GEOTX_REQUIRE(num_corr > 0 && num_corr <= SPM_MAX_K, "geotx_superpoint_matching: num_corr %d not in [1, %d]", num_corr,
                  SPM_MAX_K);
    GEOTX_REQUIRE(((uintptr_t)workspace & 255) == 0, "geotx_superpoint_matching: workspace must be 256-byte aligned");
    GEOTX_REQUIRE((((uintptr_t)ref_feats | (uintptr_t)src_feats | (uintptr_t)ref_unit_out | (uintptr_t)src_unit_out |
                    (uintptr_t)score_matrix) & 15) == 0,
                  "geotx_superpoint_matching: feature / matrix pointers must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    SpmWorkspace w = spm_layout(workspace, n, m, d, num_corr);
    float* ref_unit = ref_unit_out ? ref_unit_out : w.ref_unit;
    float* src_unit = src_unit_out ? src_unit_out : w.src_unit;
    float* matrix = score_matrix ? score_matrix : w.scores;
    const int tiles_n = (n + SPM_TILE - 1) / SPM_TILE, tiles_m = (m + SPM_TILE - 1) / SPM_TILE;
    const int prep_blocks = normalize ? (n + m + 7) / 8 : 1;
    spm_prepare_kernel<<<prep_blocks, 256, 0, st>>>(ref_feats, src_feats, n, m, d, normalize, ref_unit, src_unit, w.state,
                                                    w.hist);
    const float* a = normalize ? ref_unit : ref_feats;
    const float* b = normalize ? src_unit : src_feats;
    spm_scores_kernel<<<dim3(tiles_m, tiles_n), 256, 0, st>>>(a, b, n, m, d, ref_keep, src_keep, dual_normalization,
                                                              matrix, w.row_part, w.col_part);
    spm_sums_kernel<<<(n + m + 255) / 256, 256, 0, st>>>(w.row_part, w.col_part, n, m, tiles_n, tiles_m,
                                                         dual_normalization, ref_keep, src_keep, num_corr, w.row_sum,
                                                         w.col_sum, w.state);
    const size_t total = (size_t)n * m;
    int scan_blocks = (int)((total + 256 * 8 - 1) / (256 * 8));  // ~8 entries per thread
    if (scan_blocks > 148 * 8) scan_blocks = 148 * 8;
    if (scan_blocks < 1) scan_blocks = 1;
    spm_pass_kernel<true><<<scan_blocks, 256, 0, st>>>(matrix, n, m, dual_normalization, w.row_sum, w.col_sum, ref_keep,
                                                       src_keep, 0, w.state, w.hist);
    for (int pass = 1; pass < SPM_PASSES; ++pass)
        spm_pass_kernel<false><<<scan_blocks, 256, 0, st>>>(matrix, n, m, dual_normalization, w.row_sum, w.col_sum,
                                                            ref_keep, src_keep, pass, w.state, w.hist);
    int sort_size = 1;
    while (sort_size < num_corr) sort_size <<= 1;
    spm_collect_kernel<<<scan_blocks, 256, sizeof(unsigned long long) * sort_size, st>>>(
        matrix, n, m, w.state, w.list, num_corr, ref_idx_out, src_idx_out, score_out, count_out);
    return launch_status("geotx_superpoint_matching");
}
