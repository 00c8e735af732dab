// K7, specialised kernel for the shipped patch sizes (included by k7_log_optimal_transport.cu).
//
// Both dimensions of the padded matrix rounded up to a multiple of 8 equal P (72 for 64-point patches, 136 for
// 128-point patches; experiments/*/config.py:93).
//
// Stabilised scaling form of the same iteration (the log-domain updates of learnable_sinkhorn.py:16-18, evaluated
// without one exp per matrix entry per half-iteration).  With "absorbed" potentials ua, va (base-2 logs) the kernel
// keeps   Kt[i][j] = 2^(S2[i][j] + ua[i] + va[j])   on chip and the scaling vectors
//         a[i] = 2^(u[i] - ua[i]),   b[j] = 2^(v[j] - va[j]).   Then
//     logsumexp2_j(S2[i][j] + v[j]) = lg2(sum_j Kt[i][j] b[j]) - ua[i]        (row pass: FMAs only)
//     logsumexp2_i(S2[i][j] + u[i]) = lg2(sum_i Kt[i][j] a[i]) - va[j]        (column pass: FMAs only)
// and u = log_mu - lse, v = log_nu - lse exactly as in the reference; u and v themselves stay in the log domain, so
// nothing can overflow there.  Whenever a new potential has moved more than 24 (a factor 2^24) away from its absorbed
// value, the whole CTA re-absorbs: ua = u, va = v, Kt rebuilt from the scores (one ex2 per entry), a = b = 1.  Kt
// entries are <= 1 at every absorption (initially ua = -rowmax; later the rows or columns of Kt sum to mu or nu), so
// all sums stay within 2^(+-31) of 1: fp32 never overflows, and entries flushed to zero are < 2^-70 of their sum.
// In practice the matrix is rebuilt once or twice in the first iterations and the remaining ~98 iterations are two
// FMA passes each.  Masked rows / columns (log_mu or log_nu = -inf_value) have Kt = 0 and the
// reference's own result for them, u = 0 / v = 0 (-1e12 - (-1e12 + log n) in fp32).
//
// The potentials are carried as (absorbed value, scaling): u = ua + lg2(a) is only formed at an absorption and at the
// end, because the update needs no logarithm at all: a_new = 2^(u_new - ua) = 2^log_mu / sum (one reciprocal).
//
// Layout: Kt lives in REGISTERS.  P/8 warps; a lane owns the same 2 x P/8 float2 of the matrix in every iteration:
// for the row pass the float2 columns l + 4t (l = lane & 3) of row 8 w + (lane >> 2), for the column pass the rows
// (lane >> 2) + 8t of the two columns 8 w + 2 (lane & 3).  Shared memory only holds the vectors (scalings, potentials);
// a pass is P/8 broadcast LDS.64 + P/8 packed FFMA2 per lane, a 4- or 8-lane shuffle sum, one reciprocal, one barrier.
#pragma once

namespace geotx {

constexpr float LOT_FAR_HI = 16777216.0f;        // 2^24: re-absorb when a scaling leaves [2^-24, 2^24]
constexpr float LOT_FAR_LO = 1.0f / 16777216.0f;

__device__ __forceinline__ float lot_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

template <int P>
__global__ void __launch_bounds__(P * 4, (P == 72) ? 2 : 1)
log_optimal_transport_tile_kernel(const float* __restrict__ scores, const uint8_t* __restrict__ row_masks,
                                  const uint8_t* __restrict__ col_masks, const float* __restrict__ alpha_p,
                                  int iterations, float inf, int M, int N, float* __restrict__ out) {
    constexpr int STEPS = P / 8;
    constexpr int NW = P / 8;
    constexpr bool PRELOAD = (P == 72);  // P = 136 has no registers to spare (96 per thread: 5 warps on one SM quarter)
    __shared__ __align__(16) float ua[P];   // absorbed potentials (base-2 logs); 0 on masked / padding lines
    __shared__ __align__(16) float va[P];
    __shared__ __align__(16) float a[P];    // scalings 2^(u - ua), 2^(v - va); 1 on masked / padding lines
    __shared__ __align__(16) float bsc[P];
    __shared__ __align__(16) float log_mu[P];
    __shared__ __align__(16) float log_nu[P];
    __shared__ int valid_s[2];
    __shared__ int far_flag;
    const int rows = M + 1, cols = N + 1;
    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float alpha = __ldg(alpha_p);
    const uint8_t* rm = row_masks ? row_masks + (size_t)b * M : nullptr;  // nonzero = valid
    const uint8_t* cm = col_masks ? col_masks + (size_t)b * N : nullptr;
    const float* sb = scores + (size_t)b * M * N;
    // learnable_sinkhorn.py:45-48: the padded, masked score of entry (i, j), i < rows, j < cols (natural-log domain)
    auto padded_score = [&](int i, int j) -> float {
        const bool masked = (i < M && rm && rm[i] == 0) || (j < N && cm && cm[j] == 0);
        const float val = (i < M && j < N) ? __ldg(sb + (size_t)i * N + j) : alpha;
        return masked ? -inf : val;
    };

    if (tid < 2) valid_s[tid] = 0;
    if (tid == 0) far_flag = -1;
    __syncthreads();
    int cr = 0, cc = 0;  // :50-51
    for (int i = tid; i < M; i += NW * 32) cr += rm ? (rm[i] != 0) : 1;
    for (int j = tid; j < N; j += NW * 32) cc += cm ? (cm[j] != 0) : 1;
    cr = __reduce_add_sync(0xffffffffu, cr);
    cc = __reduce_add_sync(0xffffffffu, cc);
    if (lane == 0) { atomicAdd(&valid_s[0], cr); atomicAdd(&valid_s[1], cc); }
    __syncthreads();
    const float nvr = (float)valid_s[0], nvc = (float)valid_s[1];
    const float norm = -logf(__fadd_rn(nvr, nvc));  // :52
    for (int i = tid; i < P; i += NW * 32) {  // :54-62, :15
        float val = i < M ? norm : __fadd_rn(logf(nvc), norm);
        if (i < M && rm && rm[i] == 0) val = -inf;
        log_mu[i] = val * LOT_LOG2E;
        float w = i < N ? norm : __fadd_rn(logf(nvr), norm);
        if (i < N && cm && cm[i] == 0) w = -inf;
        log_nu[i] = w * LOT_LOG2E;
        va[i] = 0.f;
        a[i] = 1.f;
        bsc[i] = 1.f;
    }
    // first absorption: ua[i] = -max_j S2[i][j] (0 for masked and padding rows, whose Kt is 0 anyway)
    for (int i = wid; i < P; i += NW) {
        float m = -__int_as_float(0x7f800000);
        if (i < rows)
            for (int j = lane; j < cols; j += 32) m = fmaxf(m, padded_score(i, j));
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
        const bool live = i < rows && m > -0.5f * inf;
        if (lane == 0) ua[i] = (live && iterations > 0) ? -m * LOT_LOG2E : 0.f;
    }
    __syncthreads();

    const int q4 = lane & 3, q8 = lane >> 2;
    const int i = wid * 8 + q8;       // this lane's row in the row pass
    const int j0 = wid * 8 + 2 * q4;  // this lane's two columns in the column pass
    const float lmu = log_mu[i];
    const float2 lnu = *reinterpret_cast<const float2*>(log_nu + j0);
    // lines that take part: real (not padding) and not masked; the others keep u = 0 / v = 0
    const bool ract = i < rows && lmu > -0.5f * inf * LOT_LOG2E;
    const bool cact0 = j0 < cols && lnu.x > -0.5f * inf * LOT_LOG2E;
    const bool cact1 = j0 + 1 < cols && lnu.y > -0.5f * inf * LOT_LOG2E;
    const float mu2 = lot_ex2(lmu);                       // 2^log_mu, 2^log_nu: the marginals themselves
    const float2 nu2 = make_float2(lot_ex2(lnu.x), lot_ex2(lnu.y));
    const float2* b2 = reinterpret_cast<const float2*>(bsc) + q4;
    const float* acol = a + q8;

    float2 kr[STEPS], kc[STEPS];  // this lane's entries of Kt for the row pass and for the column pass
    // Kt = 2^(S2 + ua + va) from the scores (one ex2 per entry and copy; padding entries are 0)
    auto kt = [&](int r, int c) -> float {
        if (r >= rows || c >= cols) return 0.f;
        return lot_ex2(__fmaf_rn(padded_score(r, c), LOT_LOG2E, ua[r] + va[c]));
    };
    auto absorb = [&]() {
#pragma unroll
        for (int t = 0; t < STEPS; ++t) {
            const int c = 2 * (q4 + 4 * t);
            kr[t] = make_float2(kt(i, c), kt(i, c + 1));
            kc[t] = make_float2(kt(q8 + 8 * t, j0), kt(q8 + 8 * t, j0 + 1));
        }
    };
    // fold the scalings into the absorbed potentials and rebuild Kt (all threads; barriers inside)
    auto reabsorb = [&]() {
        __syncthreads();  // every thread has read the flag and its vector values
        for (int t = tid; t < P; t += NW * 32) {
            ua[t] += lot_lg2(fminf(fmaxf(a[t], 7.9e-31f), 1.2e30f));
            va[t] += lot_lg2(fminf(fmaxf(bsc[t], 7.9e-31f), 1.2e30f));
        }
        __syncthreads();
        for (int t = tid; t < P; t += NW * 32) { a[t] = 1.f; bsc[t] = 1.f; }
        absorb();
        __syncthreads();
    };
    absorb();

    // A lane that sees a scaling leave [2^-24, 2^24] stamps `far_flag` with the number of the pass; everybody reads
    // the flag together with the vector of the next pass (no reduction barrier, nothing to reset).
    int stamp = 0;
    for (int it = 0; it < iterations; ++it) {
        {  // :17  u = log_mu - logsumexp(scores + v, dim=2)   <=>   a = mu / (Kt b)
            float2 bv[STEPS];  // P = 72: issued before the flag is looked at (one shared-memory latency, not two)
            if (PRELOAD) {
#pragma unroll
                for (int t = 0; t < STEPS; ++t) bv[t] = b2[4 * t];
            }
            const bool far = far_flag == stamp;
            if (far) reabsorb();
            if (!PRELOAD || far) {
#pragma unroll
                for (int t = 0; t < STEPS; ++t) bv[t] = b2[4 * t];
            }
            ++stamp;
            float2 acc0 = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
#pragma unroll
            for (int t = 0; t < STEPS; ++t) {
                if (t & 1) acc1 = __ffma2_rn(kr[t], bv[t], acc1);
                else acc0 = __ffma2_rn(kr[t], bv[t], acc0);
            }
            const float2 s2 = __fadd2_rn(acc0, acc1);
            const float an = mu2 * lot_rcp(group_sum<4>(s2.x + s2.y));
            if (ract && q4 == 0) a[i] = an;
            if (ract && !(an > LOT_FAR_LO && an < LOT_FAR_HI)) far_flag = stamp;
        }
        __syncthreads();
        {  // :18  v = log_nu - logsumexp(scores + u, dim=1)   <=>   b = nu / (Kt^T a)
            float av[STEPS];
            if (PRELOAD) {
#pragma unroll
                for (int t = 0; t < STEPS; ++t) av[t] = acol[8 * t];
            }
            const bool far = far_flag == stamp;
            if (far) reabsorb();
            if (!PRELOAD || far) {
#pragma unroll
                for (int t = 0; t < STEPS; ++t) av[t] = acol[8 * t];
            }
            ++stamp;
            float2 acc0 = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
#pragma unroll
            for (int t = 0; t < STEPS; ++t) {
                if (t & 1) acc1 = __ffma2_rn(kc[t], make_float2(av[t], av[t]), acc1);
                else acc0 = __ffma2_rn(kc[t], make_float2(av[t], av[t]), acc0);
            }
            const float2 s2 = __fadd2_rn(acc0, acc1);
            const float bx = nu2.x * lot_rcp(group_sum<8>(s2.x)), by = nu2.y * lot_rcp(group_sum<8>(s2.y));
            if (q8 == 0) {
                if (cact0) bsc[j0] = bx;
                if (cact1) bsc[j0 + 1] = by;
            }
            if ((cact0 && !(bx > LOT_FAR_LO && bx < LOT_FAR_HI)) || (cact1 && !(by > LOT_FAR_LO && by < LOT_FAR_HI)))
                far_flag = stamp;
        }
        __syncthreads();
    }
    if (far_flag == stamp) reabsorb();  // keeps lg2 of the scalings below in its accurate range
    float* ob = out + (size_t)b * rows * cols;  // :19, :65 on the unscaled scores; u = ua + lg2(a), v = va + lg2(b)
    for (int r = wid; r < rows; r += NW) {
        const float ui = (ua[r] + lot_lg2(a[r])) * LOT_LN2;
        for (int j = lane; j < cols; j += 32) {
            const float vj = (va[j] + lot_lg2(bsc[j])) * LOT_LN2;
            ob[(size_t)r * cols + j] = __fsub_rn(__fadd_rn(__fadd_rn(padded_score(r, j), ui), vj), norm);
        }
    }
}

}  // namespace geotx
