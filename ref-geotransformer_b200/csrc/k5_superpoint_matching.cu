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
#include <stdlib.h>

namespace geotx {

constexpr int SPM_TILE = 64;        // scores tile edge
#ifndef SPM_SMALL_KC
#define SPM_SMALL_KC 32  // contraction chunk of the 64 x 64 tile (16 for the 128 x 128 one)
#endif
#ifndef SPM_SCAN_MAX_CTAS
#define SPM_SCAN_MAX_CTAS (148 * 8)  // histogram flushes cost CTAs x bins global atomics
#endif
#ifndef SPM_BIG_TILE_MIN_CTAS
#define SPM_BIG_TILE_MIN_CTAS 256
#endif
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
// spm_scores: (16 TM) x (16 TN) tile per CTA, 256 threads, TM x TN outputs per thread (rows ty*4 + 64 q, columns
// tx*4 + 64 q: every shared-memory read is a conflict-free or broadcast LDS.128), k chunks of 16 staged transposed in
// shared memory with a register prefetch of the next chunk.  Dot products accumulate in ascending k with FMA.
// ---------------------------------------------------------------------------------------------------------------
template <int TM, int TN, int KC>
__global__ void __launch_bounds__(256)
spm_scores_kernel(const float* __restrict__ a, const float* __restrict__ b, int n, int m, int d,
                  const uint8_t* __restrict__ ref_keep, const uint8_t* __restrict__ src_keep, int dual,
                  float* __restrict__ scores, float* __restrict__ row_part, float* __restrict__ col_part) {
    constexpr int RM = 16 * TM, RN = 16 * TN;            // tile rows / columns
    constexpr int PA = RM + 4, PB = RN + 4;               // staged pitches (16-byte aligned)
    constexpr int QM = TM / 4, QN = TN / 4;               // float4 groups per thread
    constexpr int KQ = KC / 4;                            // float4 per staged row
    constexpr int LA = RM * KQ / 256, LB = RN * KQ / 256; // float4 each thread stages per chunk
    constexpr int SMEM_STAGE = 2 * KC * (PA + PB);        // floats
    constexpr int SMEM_COL = 16 * RN;
    static_assert(RM * KQ % 256 == 0 && RN * KQ % 256 == 0, "staging must divide over 256 threads");
    __shared__ __align__(16) float s_mem[SMEM_STAGE > SMEM_COL ? SMEM_STAGE : SMEM_COL];
    float (*As)[KC][PA] = reinterpret_cast<float (*)[KC][PA]>(s_mem);
    float (*Bs)[KC][PB] = reinterpret_cast<float (*)[KC][PB]>(s_mem + 2 * KC * PA);
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int i0 = blockIdx.y * RM, j0 = blockIdx.x * RN;
    const int lrow = tid / KQ, lk = (tid % KQ) * 4;  // staging: rows lrow + (256 / KQ) u, k offsets lk..lk+3
    constexpr int LSTEP = 256 / KQ;
    const bool vec = (d & 3) == 0;
    auto fetch = [&](const float* base, int row, int rows, int k) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < rows) {
            const float* p = base + (size_t)row * d;
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
    float4 ra[LA], rb[LB];
    auto fetch_all = [&](int k) {
#pragma unroll
        for (int u = 0; u < LA; ++u) ra[u] = fetch(a, i0 + lrow + LSTEP * u, n, k + lk);
#pragma unroll
        for (int u = 0; u < LB; ++u) rb[u] = fetch(b, j0 + lrow + LSTEP * u, m, k + lk);
    };
    auto stage_all = [&](int buf) {
#pragma unroll
        for (int u = 0; u < LA; ++u) {
            As[buf][lk][lrow + LSTEP * u] = ra[u].x;
            As[buf][lk + 1][lrow + LSTEP * u] = ra[u].y;
            As[buf][lk + 2][lrow + LSTEP * u] = ra[u].z;
            As[buf][lk + 3][lrow + LSTEP * u] = ra[u].w;
        }
#pragma unroll
        for (int u = 0; u < LB; ++u) {
            Bs[buf][lk][lrow + LSTEP * u] = rb[u].x;
            Bs[buf][lk + 1][lrow + LSTEP * u] = rb[u].y;
            Bs[buf][lk + 2][lrow + LSTEP * u] = rb[u].z;
            Bs[buf][lk + 3][lrow + LSTEP * u] = rb[u].w;
        }
    };
    float acc[TM][TN];
#pragma unroll
    for (int r = 0; r < TM; ++r)
#pragma unroll
        for (int c = 0; c < TN; ++c) acc[r][c] = 0.f;
    const int chunks = (d + KC - 1) / KC;
    fetch_all(0);
    stage_all(0);
    __syncthreads();
    for (int ch = 0; ch < chunks; ++ch) {
        const int cur = ch & 1;
        if (ch + 1 < chunks) fetch_all((ch + 1) * KC);
#pragma unroll
        for (int kk = 0; kk < KC; ++kk) {
            float ar[TM], br[TN];
#pragma unroll
            for (int q = 0; q < QM; ++q) {
                const float4 v = *(const float4*)&As[cur][kk][q * 64 + ty * 4];
                ar[4 * q] = v.x, ar[4 * q + 1] = v.y, ar[4 * q + 2] = v.z, ar[4 * q + 3] = v.w;
            }
#pragma unroll
            for (int q = 0; q < QN; ++q) {
                const float4 v = *(const float4*)&Bs[cur][kk][q * 64 + tx * 4];
                br[4 * q] = v.x, br[4 * q + 1] = v.y, br[4 * q + 2] = v.z, br[4 * q + 3] = v.w;
            }
#pragma unroll
            for (int r = 0; r < TM; ++r)
#pragma unroll
                for (int c = 0; c < TN; ++c) acc[r][c] = fmaf(ar[r], br[c], acc[r][c]);
        }
        if (ch + 1 < chunks) stage_all(cur ^ 1);
        __syncthreads();
    }
    // exp(-clamp(2 - 2 xy, 0)); removed rows / columns contribute 0 to the sums (superpoint_matching.py:32-37)
    auto row_of = [&](int r) { return i0 + (r >> 2) * 64 + ty * 4 + (r & 3); };
    auto col_of = [&](int c) { return j0 + (c >> 2) * 64 + tx * 4 + (c & 3); };
    bool keep_c[TN];
#pragma unroll
    for (int c = 0; c < TN; ++c) {
        const int j = col_of(c);
        keep_c[c] = j < m && (!src_keep || src_keep[j]);
    }
    float cs[TN];
#pragma unroll
    for (int c = 0; c < TN; ++c) cs[c] = 0.f;
#pragma unroll
    for (int r = 0; r < TM; ++r) {
        const int i = row_of(r);
        const bool keep_r = i < n && (!ref_keep || ref_keep[i]);
        float rs = 0.f;
#pragma unroll
        for (int c = 0; c < TN; ++c) {
            const float d2 = fmaxf(2.0f - 2.0f * acc[r][c], 0.0f);
            const float s = (keep_r && keep_c[c]) ? expf(-d2) : 0.f;
            acc[r][c] = s;
            rs += s;
            cs[c] += s;
        }
        if (i < n) {
#pragma unroll
            for (int q = 0; q < QN; ++q) {
                const int j = col_of(4 * q);
                float* out = scores + (size_t)i * m + j;
                if ((m & 3) == 0 && j + 3 < m) {
                    *(float4*)out = make_float4(acc[r][4 * q], acc[r][4 * q + 1], acc[r][4 * q + 2], acc[r][4 * q + 3]);
                } else {
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        if (j + c < m) out[c] = acc[r][4 * q + c];
                }
            }
        }
        if (dual) {  // row partial: the 16 threads of a row group are 16 consecutive lanes -> fixed xor tree
#pragma unroll
            for (int off = 8; off > 0; off >>= 1) rs += __shfl_xor_sync(0xffffffffu, rs, off);
            if (tx == 0 && i < n) row_part[(size_t)blockIdx.x * n + i] = rs;
        }
    }
    if (!dual) return;
    // column partials: through shared memory (the staging buffers are dead), added in ascending ty
    float (*s_col)[RN] = reinterpret_cast<float (*)[RN]>(s_mem);
#pragma unroll
    for (int c = 0; c < TN; ++c) s_col[ty][col_of(c) - j0] = cs[c];
    __syncthreads();
    for (int t = tid; t < RN; t += 256) {
        if (j0 + t < m) {
            float sum = 0.f;
#pragma unroll
            for (int y = 0; y < 16; ++y) sum += s_col[y][t];
            col_part[(size_t)blockIdx.y * m + j0 + t] = sum;
        }
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

// Last CTA of a pass: find the digit at which the histogram, summed from the top, first covers `need` keys.
// 256 threads sum 8 bins each (thread 0 = bins 2047..2040); warp 0 scans 8 of those per lane with shuffles, and the
// one lane in whose range the count crosses `need` walks its <= 64 bins.
__device__ void spm_fix_digit(SpmState* state, const unsigned int* hist, int pass, unsigned int* s_bins) {
    __shared__ unsigned int s_part[256];
    const int tid = threadIdx.x;
    for (int i = tid; i < SPM_BINS; i += blockDim.x) s_bins[i] = __ldcg(hist + i);  // the walk below must not chase L2
    __syncthreads();
    unsigned int local = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) local += s_bins[SPM_BINS - 1 - (tid * 8 + b)];
    s_part[tid] = local;
    __syncthreads();
    if (tid >= 32) return;
    const unsigned int need = state->need;
    unsigned int mine = 0;
#pragma unroll
    for (int g = 0; g < 8; ++g) mine += s_part[tid * 8 + g];
    unsigned int incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const unsigned int up = __shfl_up_sync(0xffffffffu, incl, off);
        if (tid >= off) incl += up;
    }
    const unsigned int covers = __ballot_sync(0xffffffffu, incl >= need);
    const int owner = covers ? __ffs(covers) - 1 : 31;  // no lane covers `need` only if K_eff exceeded the candidates
    if (tid != owner) return;
    unsigned int cum = incl - mine;
    int bin = SPM_BINS - 1 - tid * 64;
    const int last = bin - 63;
    unsigned int h = s_bins[bin];
    while (bin > last && cum + h < need) {
        cum += h;
        --bin;
        h = s_bins[bin];
    }
    state->prefix |= (unsigned long long)bin << c_spm_shift[pass];
    if (cum + h <= need) {  // the bucket is taken whole: keys >= prefix are exactly the selection
        state->resolved = 1u;
    } else {
        state->need = need - cum;
    }
}

// Histogram increment of a whole warp.  The scores are strongly clustered (nearly every entry shares its top digits),
// and 32 shared-memory atomics on one bin serialise: when all lanes agree, one lane adds 32; otherwise the digits are
// spread and per-lane atomics rarely collide.  digit == ~0u: nothing to add.
__device__ __forceinline__ void spm_hist_add(unsigned int* s_hist, unsigned int digit, int lane) {
    const unsigned int first = __shfl_sync(0xffffffffu, digit, 0);
    if (__all_sync(0xffffffffu, digit == first)) {
        if (lane == 0 && digit != ~0u) atomicAdd(&s_hist[digit], 32u);
    } else if (digit != ~0u) {
        atomicAdd(&s_hist[digit], 1u);
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
    const int lane = threadIdx.x & 31;
    const int shift = c_spm_shift[pass], width = c_spm_width[pass];
    const unsigned long long prefix = FIRST ? 0ull : state->prefix;
    const int above = shift + width;  // bits above this digit must equal the prefix
    const unsigned int dmask = (1u << width) - 1u;
    const size_t total = (size_t)n * m;
    // entry idx -> its digit (or ~0u); FIRST also turns the raw score into the final one, in place
    auto visit = [&](uint32_t raw, size_t idx, int i, int j) -> uint32_t {
        uint32_t bits = raw;
        if (FIRST) {
            const bool keep = (!ref_keep || ref_keep[i]) && (!src_keep || src_keep[j]);
            float f = __uint_as_float(raw);
            if (dual) f = __fdiv_rn(f, __ldg(row_sum + i)) * __fdiv_rn(f, __ldg(col_sum + j));  // superpoint_matching.py:39-41
            bits = keep ? __float_as_uint(f) : SPM_MASKED;
        }
        return bits;
    };
    auto digit_of = [&](uint32_t bits, size_t idx) -> unsigned int {
        if (bits == SPM_MASKED) return ~0u;
        const unsigned long long key = spm_key(bits, (uint32_t)idx);
        if (!FIRST && (key >> above) != (prefix >> above)) return ~0u;
        return (unsigned int)(key >> shift) & dmask;
    };
    const size_t quads = total >> 2;
    uint4* q4 = (uint4*)scores;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q - lane < quads; q += stride) {  // warp-uniform bound
        const bool live = q < quads;
        uint4 v = make_uint4(SPM_MASKED, SPM_MASKED, SPM_MASKED, SPM_MASKED);
        const size_t idx = q << 2;
        if (live) {
            v = q4[q];
            if (FIRST) {
                int i = (int)(idx / (unsigned)m), j = (int)(idx - (size_t)i * m);
                uint32_t* e = &v.x;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    e[t] = visit(e[t], idx + t, i, j);
                    if (++j == m) j = 0, ++i;
                }
                q4[q] = v;
            }
        }
        spm_hist_add(s_hist, live ? digit_of(v.x, idx) : ~0u, lane);
        spm_hist_add(s_hist, live ? digit_of(v.y, idx + 1) : ~0u, lane);
        spm_hist_add(s_hist, live ? digit_of(v.z, idx + 2) : ~0u, lane);
        spm_hist_add(s_hist, live ? digit_of(v.w, idx + 3) : ~0u, lane);
    }
    if (blockIdx.x == 0 && threadIdx.x < 32) {  // the last total % 4 entries
        const size_t idx = (quads << 2) + lane;
        unsigned int digit = ~0u;
        if (lane < (int)(total & 3)) {
            uint32_t bits = ((uint32_t*)scores)[idx];
            if (FIRST) {
                const int i = (int)(idx / (unsigned)m), j = (int)(idx - (size_t)i * m);
                bits = visit(bits, idx, i, j);
                ((uint32_t*)scores)[idx] = bits;
            }
            digit = digit_of(bits, idx);
        }
        spm_hist_add(s_hist, digit, lane);
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
        auto take = [&](uint32_t bits, size_t idx) {
            if (bits == SPM_MASKED) return;
            const unsigned long long key = spm_key(bits, (uint32_t)idx);
            if (key >= threshold) {
                const unsigned int pos = atomicAdd(&state->list_count, 1u);
                if (pos < (unsigned int)num_corr) list[pos] = key;
            }
        };
        const size_t quads = total >> 2;
        const uint4* q4 = (const uint4*)scores;
        for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < quads; q += (size_t)gridDim.x * blockDim.x) {
            const uint4 v = q4[q];
            const size_t idx = q << 2;
            take(v.x, idx), take(v.y, idx + 1), take(v.z, idx + 2), take(v.w, idx + 3);
        }
        if (blockIdx.x == 0 && threadIdx.x < (total & 3)) {
            const size_t idx = (quads << 2) + threadIdx.x;
            take(((const uint32_t*)scores)[idx], idx);
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
#ifdef SPM_DBG_STOP
    const char* stop_env = getenv("GEOTX_SPM_STOP_AFTER");
    int stop_after = stop_env ? atoi(stop_env) : 100, launched = 0;
#define SPM_STEP() do { if (++launched >= stop_after) return launch_status("geotx_superpoint_matching"); } while (0)
#else
#define SPM_STEP() do { } while (0)
#endif
    SpmWorkspace w = spm_layout(workspace, n, m, d, num_corr);
    float* ref_unit = ref_unit_out ? ref_unit_out : w.ref_unit;
    float* src_unit = src_unit_out ? src_unit_out : w.src_unit;
    float* matrix = score_matrix ? score_matrix : w.scores;
    // 128 x 128 tiles (64 accumulators per thread) once they fill the 148 SMs about twice, else 64 x 64
    const bool big_tile = (long long)((n + 127) / 128) * ((m + 127) / 128) >= SPM_BIG_TILE_MIN_CTAS;
    const int tile = big_tile ? 128 : SPM_TILE;
    const int tiles_n = (n + tile - 1) / tile, tiles_m = (m + tile - 1) / tile;
    const int prep_blocks = normalize ? (n + m + 7) / 8 : 1;
    spm_prepare_kernel<<<prep_blocks, 256, 0, st>>>(ref_feats, src_feats, n, m, d, normalize, ref_unit, src_unit, w.state,
                                                    w.hist);
    SPM_STEP();
    const float* a = normalize ? ref_unit : ref_feats;
    const float* b = normalize ? src_unit : src_feats;
    if (big_tile)
        spm_scores_kernel<8, 8, 16><<<dim3(tiles_m, tiles_n), 256, 0, st>>>(a, b, n, m, d, ref_keep, src_keep, dual_normalization,
                                                                        matrix, w.row_part, w.col_part);
    else
        spm_scores_kernel<4, 4, SPM_SMALL_KC><<<dim3(tiles_m, tiles_n), 256, 0, st>>>(a, b, n, m, d, ref_keep, src_keep, dual_normalization,
                                                                        matrix, w.row_part, w.col_part);
    SPM_STEP();
    spm_sums_kernel<<<(n + m + 255) / 256, 256, 0, st>>>(w.row_part, w.col_part, n, m, tiles_n, tiles_m,
                                                         dual_normalization, ref_keep, src_keep, num_corr, w.row_sum,
                                                         w.col_sum, w.state);
    SPM_STEP();
    const size_t total = (size_t)n * m;
    int scan_blocks = (int)((total + 256 * 8 - 1) / (256 * 8));  // two 16-byte accesses per thread
    if (scan_blocks > SPM_SCAN_MAX_CTAS) scan_blocks = SPM_SCAN_MAX_CTAS;
    if (scan_blocks < 1) scan_blocks = 1;
    spm_pass_kernel<true><<<scan_blocks, 256, 0, st>>>(matrix, n, m, dual_normalization, w.row_sum, w.col_sum, ref_keep,
                                                       src_keep, 0, w.state, w.hist);
    SPM_STEP();
    for (int pass = 1; pass < SPM_PASSES; ++pass) {
        spm_pass_kernel<false><<<scan_blocks, 256, 0, st>>>(matrix, n, m, dual_normalization, w.row_sum, w.col_sum,
                                                            ref_keep, src_keep, pass, w.state, w.hist);
        SPM_STEP();
    }
    int sort_size = 1;
    while (sort_size < num_corr) sort_size <<= 1;
    spm_collect_kernel<<<scan_blocks, 256, sizeof(unsigned long long) * sort_size, st>>>(
        matrix, n, m, w.state, w.list, num_corr, ref_idx_out, src_idx_out, score_out, count_out);
    return launch_status("geotx_superpoint_matching");
}
