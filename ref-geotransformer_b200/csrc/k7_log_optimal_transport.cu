// K7: LearnableLogOptimalTransport -- SuperGlue-style log-domain Sinkhorn with a dustbin row and
// column, one CTA per problem, the whole iteration loop inside one launch.
//
// Replaces /root/reference/geotransformer/modules/sinkhorn/learnable_sinkhorn.py:13-66 (called from
// experiments/*/model.py:187-189 on (256, 64, 64) / (256, 128, 128) patch score matrices with 100 iterations): the
// reference issues two torch.logsumexp (plus broadcast adds) per iteration, about a thousand launches per pair, each a
// pass over a (B, M+1, N+1) tensor in HBM.  Here the padded matrix lives in shared memory for the whole loop and HBM
// sees the scores once (twice: the epilogue re-reads them to add u and v to the unscaled values) and the result once.
//
// Arithmetic.  The kernel is bound by the MUFU pipe (one ex2 per matrix entry per half-iteration), so everything
// around the ex2 is kept to one shared-memory load pair, two adds and one accumulate:
//   * base-2 domain: S2 = S log2(e), u2 = u log2(e), ... so that exp(x - m) is ex2(x2 - m2) with no multiply;
//     lse2 = m2 + lg2(sum);  the epilogue converts back: out = ((S + u2 ln2) + v2 ln2) - norm (learnable_sinkhorn.py:19,65)
//   * the stabiliser m of a line is NOT its maximum but the line's lse of the previous iteration (log-sum-exp is exact
//     for any m; lse_prev >= max_prev and the potentials move little between iterations), which removes the max pass,
//     its shuffle reduction and the need to keep the line in registers.  The first iteration, and any line whose sum
//     leaves [2^-100, 2^100] (never observed), takes the true maximum like torch.logsumexp does.
//   * masked entries hold -inf_value (-1e12) as in the reference and absorb the same way (never -inf / NaN).
// Result: equal to the reference to ~1e-6 on finite entries (tests: 1e-4), masked entries to the ulp.
//
// Layout: padded matrix rows x cols (= M+1, N+1) with a row stride == 4 (mod 32) words, which makes BOTH passes
// bank-conflict free with a 4-lane-per-row / 8-lane-per-column split of the warp:
//   row pass     lane = l + 4*ii handles row r0+ii, columns l, l+4, ...   bank = 4*ii + l + const
//   column pass  lane = jj + 4*rr handles column c0+jj, rows rr, rr+8, ... bank = 4*rr + jj + const
#include "geotx_common.cuh"

namespace geotx {

constexpr float LOT_LOG2E = 1.4426950408889634f;
constexpr float LOT_LN2 = 0.6931471805599453f;

__device__ __forceinline__ float lot_ex2(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float lot_lg2(float x) {
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

template <int G>
__device__ __forceinline__ float group_sum(float s) {
#pragma unroll
    for (int off = 1; off < G; off <<= 1) s += __shfl_xor_sync(0xffffffffu, s, off * (G == 4 ? 1 : 4));
    return s;
}
template <int G>
__device__ __forceinline__ float group_max(float s) {
#pragma unroll
    for (int off = 1; off < G; off <<= 1) s = fmaxf(s, __shfl_xor_sync(0xffffffffu, s, off * (G == 4 ? 1 : 4)));
    return s;
}

// Sum over one line of the matrix of 2^(line[e] + add[e] - m), computed by G cooperating lanes; lane g of the group
// takes the elements e = g + G*t.  `line` and `add` already point at the lane's first element, `lstep` / G are the
// distances to its next one, `cnt` is its element count and `cnt_max` the group's largest (warp-uniform loop bound).
template <int G>
__device__ __forceinline__ float line_sum(const float* __restrict__ line, int lstep, const float* __restrict__ add,
                                          int cnt, int cnt_max, float m) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int t = 0;
    for (; t + 4 <= cnt_max; t += 4) {
        const float a0 = line[0], a1 = line[lstep], a2 = line[2 * lstep], a3 = line[3 * lstep];
        const float b0 = add[0], b1 = add[G], b2 = add[2 * G], b3 = add[3 * G];
        s0 += lot_ex2(a0 + b0 - m);
        s1 += lot_ex2(a1 + b1 - m);
        s2 += lot_ex2(a2 + b2 - m);
        if (t + 3 < cnt) s3 += lot_ex2(a3 + b3 - m);  // only the last element of a lane can be missing
        line += 4 * lstep;
        add += 4 * G;
    }
    for (; t < cnt_max; ++t) {
        if (t < cnt) s0 += lot_ex2(line[0] + add[0] - m);
        line += lstep;
        add += G;
    }
    return group_sum<G>((s0 + s1) + (s2 + s3));
}

template <int G>
__device__ __forceinline__ float line_max(const float* __restrict__ line, int lstep, const float* __restrict__ add,
                                          int cnt) {
    float m = -__int_as_float(0x7f800000);
    for (int t = 0; t < cnt; ++t) m = fmaxf(m, line[t * lstep] + add[t * G]);
    return group_max<G>(m);
}

// lse2 of a line: stabilised by `guess` (the previous iteration's value) when `use_guess`, by the true maximum otherwise
// or when the guess turned out to be far off.  All lanes of the warp call it together (`act`: the lane's line exists).
template <int G>
__device__ __forceinline__ float line_lse2(const float* __restrict__ line, int lstep, const float* __restrict__ add,
                                           int cnt, int cnt_max, float guess, bool use_guess, bool act) {
    float m = guess, s = 1.f;
    if (use_guess) s = line_sum<G>(line, lstep, add, cnt, cnt_max, m);
    const bool redo = act && !(use_guess && s > 7.9e-31f && s < 1.2e30f);
    if (__any_sync(0xffffffffu, redo)) {  // warp-uniform: the shuffles inside need every lane
        const float mt = line_max<G>(line, lstep, add, cnt);
        const float st = line_sum<G>(line, lstep, add, cnt, cnt_max, mt);
        if (redo) { m = mt; s = st; }
    }
    return m + lot_lg2(s);
}

constexpr int LOT_MAX_WARPS = 20;

__global__ void __launch_bounds__(LOT_MAX_WARPS * 32, 2)
log_optimal_transport_kernel(const float* __restrict__ scores, const uint8_t* __restrict__ row_masks,
                             const uint8_t* __restrict__ col_masks, const float* __restrict__ alpha_p, int iterations,
                             float inf, int M, int N, int stride, float* __restrict__ out) {
    extern __shared__ float smem[];
    const int rows = M + 1, cols = N + 1;
    float* S = smem;                    // rows x stride, base-2 domain
    float* u = S + (size_t)rows * stride;
    float* v = u + rows;
    float* log_mu = v + cols;
    float* log_nu = log_mu + rows;
    float* lse_r = log_nu + cols;       // previous lse2 of every row / column: the next stabiliser
    float* lse_c = lse_r + rows;
    __shared__ int valid_s[2];
    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarps = blockDim.x >> 5;
    const float alpha = __ldg(alpha_p);
    const uint8_t* rm = row_masks ? row_masks + (size_t)b * M : nullptr;  // nonzero = valid
    const uint8_t* cm = col_masks ? col_masks + (size_t)b * N : nullptr;

    if (tid < 2) valid_s[tid] = 0;
    __syncthreads();
    // learnable_sinkhorn.py:50-51: num_valid_row / num_valid_col
    int cr = 0, cc = 0;
    for (int i = tid; i < M; i += blockDim.x) cr += rm ? (rm[i] != 0) : 1;
    for (int j = tid; j < N; j += blockDim.x) cc += cm ? (cm[j] != 0) : 1;
    cr = __reduce_add_sync(0xffffffffu, cr);
    cc = __reduce_add_sync(0xffffffffu, cc);
    if (lane == 0) { atomicAdd(&valid_s[0], cr); atomicAdd(&valid_s[1], cc); }
    // :45-48 padded scores: [scores | alpha] over [alpha ... alpha], masked rows / columns filled with -inf
    const float* sb = scores + (size_t)b * M * N;
    for (int i = wid; i < rows; i += nwarps) {
        const bool rmask = i < M && rm && rm[i] == 0;
        for (int j = lane; j < cols; j += 32) {
            const bool cmask = j < N && cm && cm[j] == 0;
            float val = (i < M && j < N) ? __ldg(sb + (size_t)i * N + j) : alpha;
            if (rmask || cmask) val = -inf;
            S[i * stride + j] = val * LOT_LOG2E;
        }
    }
    __syncthreads();
    const float nvr = (float)valid_s[0], nvc = (float)valid_s[1];
    const float norm = -logf(__fadd_rn(nvr, nvc));  // :52
    // :54-62 log_mu / log_nu; u = v = 0 (:15)
    for (int i = tid; i < rows; i += blockDim.x) {
        float val = i < M ? norm : __fadd_rn(logf(nvc), norm);
        if (i < M && rm && rm[i] == 0) val = -inf;
        log_mu[i] = val * LOT_LOG2E;
        u[i] = 0.f;
    }
    for (int j = tid; j < cols; j += blockDim.x) {
        float val = j < N ? norm : __fadd_rn(logf(nvr), norm);
        if (j < N && cm && cm[j] == 0) val = -inf;
        log_nu[j] = val * LOT_LOG2E;
        v[j] = 0.f;
    }
    __syncthreads();

    // row pass: 8 rows per warp, 4 lanes each; column pass: 4 columns per warp, 8 lanes each.  A group whose line does
    // not exist works on line 0 and discards the result, so that control flow stays warp-uniform.
    const int l4 = lane & 3, i8 = lane >> 2;
    const int j4 = lane & 3, r8 = lane >> 2;
    const int rcnt = (cols - l4 + 3) >> 2, rcnt_max = (cols + 3) >> 2;  // elements of this lane in a row / a column
    const int ccnt = (rows - r8 + 7) >> 3, ccnt_max = (rows + 7) >> 3;
    for (int it = 0; it < iterations; ++it) {
        // :17  u = log_mu - logsumexp(scores + v, dim=2)
        for (int r0 = wid * 8; r0 < rows; r0 += nwarps * 8) {
            const int i = r0 + i8;
            const bool act = i < rows;
            const int ii = act ? i : 0;
            const float lse = line_lse2<4>(S + (size_t)ii * stride + l4, 4, v + l4, rcnt, rcnt_max, lse_r[ii], it > 0, act);
            if (act && l4 == 0) { u[i] = log_mu[i] - lse; lse_r[i] = lse; }
        }
        __syncthreads();
        // :18  v = log_nu - logsumexp(scores + u, dim=1)
        for (int c0 = wid * 4; c0 < cols; c0 += nwarps * 4) {
            const int j = c0 + j4;
            const bool act = j < cols;
            const int jj = act ? j : 0;
            const float lse = line_lse2<8>(S + (size_t)r8 * stride + jj, 8 * stride, u + r8, ccnt, ccnt_max, lse_c[jj],
                                           it > 0, act);
            if (act && r8 == 0) { v[j] = log_nu[j] - lse; lse_c[j] = lse; }
        }
        __syncthreads();
    }
    // :19, :65  (scores + u + v) - norm, on the unscaled scores
    float* ob = out + (size_t)b * rows * cols;
    for (int i = wid; i < rows; i += nwarps) {
        const float ui = u[i] * LOT_LN2;
        const bool rmask = i < M && rm && rm[i] == 0;
        for (int j = lane; j < cols; j += 32) {
            const bool cmask = j < N && cm && cm[j] == 0;
            float val = (i < M && j < N) ? __ldg(sb + (size_t)i * N + j) : alpha;
            if (rmask || cmask) val = -inf;
            ob[(size_t)i * cols + j] = __fsub_rn(__fadd_rn(__fadd_rn(val, ui), v[j] * LOT_LN2), norm);
        }
    }
}

}  // namespace geotx

#include "k7_tile.cuh"

// ---- C ABI --------------------------------------------------------------------------------------------------
#include "host_util.h"
using namespace geotx;

extern "C" int geotx_log_optimal_transport(const float* scores, const uint8_t* row_masks, const uint8_t* col_masks,
                                           const float* alpha, int num_iterations, float inf, int B, int M, int N,
                                           float* out, void* stream) {
    GEOTX_REQUIRE(scores && alpha && out, "geotx_log_optimal_transport: null pointer");
    GEOTX_REQUIRE(B > 0 && M > 0 && N > 0, "geotx_log_optimal_transport: bad sizes B=%d M=%d N=%d", B, M, N);
    GEOTX_REQUIRE(num_iterations >= 0, "geotx_log_optimal_transport: num_iterations=%d", num_iterations);
    const int rows = M + 1, cols = N + 1;
    cudaStream_t st = (cudaStream_t)stream;
    const int rp = (rows + 7) & ~7, cp = (cols + 7) & ~7;
    if (rp == cp && (rp == 72 || rp == 136)) {  // the shipped patch sizes (64 / 128 points): unrolled float2 kernel
        if (rp == 72)
            log_optimal_transport_tile_kernel<72><<<B, 72 * 4, 0, st>>>(scores, row_masks, col_masks, alpha,
                                                                        num_iterations, inf, M, N, out);
        else
            log_optimal_transport_tile_kernel<136><<<B, 136 * 4, 0, st>>>(scores, row_masks, col_masks, alpha,
                                                                          num_iterations, inf, M, N, out);
        return launch_status("geotx_log_optimal_transport");
    }
    int stride = cols;
    while ((stride & 31) != 4) ++stride;
    const size_t smem = ((size_t)rows * stride + 3 * (size_t)rows + 3 * (size_t)cols) * sizeof(float);
    GEOTX_REQUIRE(smem <= 227 * 1024, "geotx_log_optimal_transport: a (%d x %d) problem needs %zu bytes of shared "
                  "memory (> 227 KB)", rows, cols, smem);
    int rc = check_cuda(cudaFuncSetAttribute(log_optimal_transport_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem), "geotx_log_optimal_transport: shared memory");
    if (rc) return rc;
    int warps = max((rows + 7) / 8, (cols + 3) / 4);
    warps = min(max(warps, 4), LOT_MAX_WARPS);  // more line groups than warps: the pass loops
    log_optimal_transport_kernel<<<B, warps * 32, smem, st>>>(
        scores, row_masks, col_masks, alpha, num_iterations, inf, M, N, stride, out);
    return launch_status("geotx_log_optimal_transport");
}
