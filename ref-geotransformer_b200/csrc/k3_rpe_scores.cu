// K3a: relative-position attention scores with proj_p folded into the query.
//
// Replaces /root/reference/geotransformer/modules/transformer/rpe_transformer.py:54-56
//     p = proj_p(embed_qk)                       # (B,N,M,C) x (C,C): 2 N M C^2 flop + a second N*M*C tensor
//     scores_p = einsum('bhnc,bhnmc->bhnm', q, p)
// by the exact identity
//     scores_p[b,h,i,j] = (W_p[h]^T q[b,h,i]) . E[b,i,j,:]  (+ q.b_p, constant over j: cancels in the softmax)
// so each layer reads E exactly once and never forms p.  The kernel is HBM-bound on E (2 n m C bytes in bf16):
// every warp streams rows E[b,i,j0..j0+7,:] with 16-byte loads (8 rows x 64 contiguous bytes per instruction)
// and contracts them against the folded query of row i on the tensor cores (mma.sync m16n8k16, 16-bit in / fp32
// accumulate, rows of the A fragment = heads).  The folded query is split into hi + lo 16-bit parts so that its
// rounding does not add to the error budget (E itself is the only 16-bit-rounded operand).
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "sm100_ptx.cuh"
#include "host_util.h"

namespace geotx {

constexpr int K3A_WARPS = 8;

// 16-bit storage formats of E: both are contracted with fp32 accumulation; F16 = IEEE half (the format the embedding
// kernel writes by default: 11-bit significand), otherwise bfloat16.
template <bool F16>
__device__ __forceinline__ void mma_16816(float* c, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
    // rows 8..15 of A are zero (a1 = a3 = 0): only H <= 8 heads are populated
    if (F16)
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
            : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
    else
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
            : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
}
// hi + lo split of two fp32 values into the 16-bit format (hi = rounded value, lo = rounded remainder)
template <bool F16>
__device__ __forceinline__ void split_hi_lo(float x, float y, uint32_t& hi, uint32_t& lo) {
    if (F16) {
        // saturate instead of overflowing to inf (|q| > 65504 does not occur with sane weights, but inf * 0 = NaN)
        asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(y), "f"(x));
        const float2 hf = __half22float2(*(const __half2*)&hi);
        asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(y - hf.y), "f"(x - hf.x));
    } else {
        const float h0 = __bfloat162float(__float2bfloat16_rn(x));
        const float h1 = __bfloat162float(__float2bfloat16_rn(y));
        hi = pack_bf16x2(h0, h1);
        lo = pack_bf16x2(x - h0, y - h1);
    }
}

__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// Streaming structure: the kernel is HBM-bound, so what matters is bytes in flight, not registers.  Every warp owns a
// private ring of K3A_STAGES shared-memory stages and is its own producer: lane 0 asks the TMA unit (cp.async.bulk) for
// the next 16 key rows of E (16 x 2C bytes, contiguous in HBM) while the warp contracts the previous ones, completion
// signalled on a per-stage mbarrier.  8 warps x 16 KB keep 128 KB per SM in flight (Little's law needs ~40).
// One bulk copy per chunk: per-row copies into a padded (conflict-free) layout were TMA-issue-bound (72 % of HBM peak);
// the unpadded layout costs a 2-way bank conflict on the fragment reads, which need <1/4 of the smem bandwidth anyway.
// Work is split into chunks of 16 keys of one query row; each warp takes a contiguous range of chunks (persistent grid,
// one CTA per SM), reloading its folded-query fragments only when the row changes.
//
// Channel permutation: lane (g = lane/4, t = lane%4) owns channels 32 s + 8 t + [0,8) of E row j0+g for every s;
// the two k16 steps of chunk s use the 8 channels as k-slots {2t,2t+1 | 2t+8,2t+9}.  Any fixed permutation of
// the contraction index is legal as long as A (the folded query) uses the same one, which it does.
// ring depth per warp, measured at n = 2048 (tools/k3_bench.py): C = 256 (8 KB stages): 2 deep 6.3 TB/s, 3 deep
// 5.8 TB/s; C = 128 (4 KB stages): 2 / 3 / 4 / 6 deep 5.75 / 5.88 / 6.11 / 5.85 TB/s -- 128 KB in flight per SM is
// the sweet spot (more outstanding bulk copies lose DRAM locality)
#ifndef K3A_ST256
#define K3A_ST256 2
#endif
#ifndef K3A_ST128
#define K3A_ST128 4
#endif
constexpr int K3A_KEYS = 16;  // keys per chunk / stage

template <int C>
struct K3ACfg {
    static constexpr int ROW_BYTES = 2 * C;
    static constexpr int PITCH = ROW_BYTES;  // one bulk copy per chunk; the 2-way bank conflict of the reads is free here
    static constexpr int STAGE_BYTES = K3A_KEYS * PITCH;
    static constexpr int STAGES = (C == 256) ? K3A_ST256 : K3A_ST128;
    static constexpr int SMEM_BYTES = K3A_WARPS * STAGES * STAGE_BYTES + K3A_WARPS * STAGES * 8 + 128;
};

template <int C, bool F16>
__global__ void __launch_bounds__(K3A_WARPS * 32, 1)
rpe_scores_tc_kernel(const float* __restrict__ qt, long long qsb, long long qsh, long long qsn,
                       const void* __restrict__ E,
                       int H, int n, int m, int row_begin, int row_count, long long chunks_total, int chunks_per_row,
                       long long chunks_per_warp, float* __restrict__ scores) {
    using Cfg = K3ACfg<C>;
    constexpr int K3A_STAGES = Cfg::STAGES;
    constexpr int S = C / 32;
    extern __shared__ __align__(128) uint8_t k3a_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    uint8_t* ring = k3a_smem + (size_t)warp * K3A_STAGES * Cfg::STAGE_BYTES;
    const uint32_t ring_u32 = smem_u32(ring);
    const uint32_t bars = smem_u32(k3a_smem + (size_t)K3A_WARPS * K3A_STAGES * Cfg::STAGE_BYTES) + warp * K3A_STAGES * 8;

    const long long wid = (long long)blockIdx.x * K3A_WARPS + warp;
    const long long c_begin = wid * chunks_per_warp;
    long long c_end = c_begin + chunks_per_warp;
    if (c_end > chunks_total) c_end = chunks_total;
    if (lane == 0) {
        for (int s = 0; s < K3A_STAGES; ++s) ptx::mbar_init(bars + 8 * s, 1);
        ptx::fence_barrier_init();
    }
    __syncwarp();
    if (c_begin >= c_end) return;

    // lane 0: ask the TMA unit for chunk c (up to 16 key rows of one query row) into stage c % K3A_STAGES
    auto issue = [&](long long c) {
        const long long r = c / chunks_per_row;
        const int j0 = (int)(c - r * chunks_per_row) * K3A_KEYS;
        const int keys = min(K3A_KEYS, m - j0);
        const uint32_t st = (uint32_t)((c - c_begin) % K3A_STAGES);
        const uint32_t dst = ring_u32 + st * Cfg::STAGE_BYTES;
        const uint8_t* src = (const uint8_t*)E + ((size_t)r * m + j0) * Cfg::ROW_BYTES;
        ptx::mbar_arrive_expect_tx(bars + 8 * st, (uint32_t)keys * Cfg::ROW_BYTES);
        ptx::bulk_g2s(dst, src, (uint32_t)keys * Cfg::ROW_BYTES, bars + 8 * st);
    };
    if (lane == 0) {
        for (int s = 0; s < K3A_STAGES; ++s)
            if (c_begin + s < c_end) issue(c_begin + s);
    }

    uint32_t ah[S][4], al[S][4];
    long long cur_row = -1;
    float* srow = nullptr;
    for (long long c = c_begin; c < c_end; ++c) {
        const long long r = c / chunks_per_row;
        const int j0 = (int)(c - r * chunks_per_row) * K3A_KEYS;
        if (r != cur_row) {
            // A fragments: folded query of (b, h = g, i), hi/lo split in E's 16-bit format
            cur_row = r;
            const int b = (int)(r / row_count);
            const int i = row_begin + (int)(r - (long long)b * row_count);
            srow = scores + (((size_t)b * H + g) * n + i) * m;
            if (g < H) {
                const float* qrow = qt + (size_t)b * qsb + (size_t)g * qsh + (size_t)i * qsn + 8 * t;
#pragma unroll
                for (int s = 0; s < S; ++s) {
                    const float4 u = __ldg((const float4*)(qrow + 32 * s));
                    const float4 v = __ldg((const float4*)(qrow + 32 * s + 4));
                    const float f[8] = {u.x, u.y, u.z, u.w, v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) split_hi_lo<F16>(f[2 * e], f[2 * e + 1], ah[s][e], al[s][e]);
                }
            } else {
#pragma unroll
                for (int s = 0; s < S; ++s)
#pragma unroll
                    for (int e = 0; e < 4; ++e) { ah[s][e] = 0u; al[s][e] = 0u; }
            }
        }
        const uint32_t idx = (uint32_t)(c - c_begin);
        const uint32_t st = idx % K3A_STAGES;
        ptx::mbar_wait(bars + 8 * st, (idx / K3A_STAGES) & 1);
        const uint32_t base = ring_u32 + st * Cfg::STAGE_BYTES + 16 * t;
        // two independent accumulator chains per key group (hi and lo parts) keep the mma pipe busy
        float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f};
        float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int s = 0; s < S; ++s) {
            uint4 e0, e1;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(e0.x), "=r"(e0.y), "=r"(e0.z), "=r"(e0.w) : "r"(base + g * Cfg::PITCH + 64 * s));
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(e1.x), "=r"(e1.y), "=r"(e1.z), "=r"(e1.w) : "r"(base + (g + 8) * Cfg::PITCH + 64 * s));
            mma_16816<F16>(c0, ah[s][0], ah[s][1], e0.x, e0.y);
            mma_16816<F16>(d0, al[s][0], al[s][1], e0.x, e0.y);
            mma_16816<F16>(c1, ah[s][0], ah[s][1], e1.x, e1.y);
            mma_16816<F16>(d1, al[s][0], al[s][1], e1.x, e1.y);
            mma_16816<F16>(c0, ah[s][2], ah[s][3], e0.z, e0.w);
            mma_16816<F16>(d0, al[s][2], al[s][3], e0.z, e0.w);
            mma_16816<F16>(c1, ah[s][2], ah[s][3], e1.z, e1.w);
            mma_16816<F16>(d1, al[s][2], al[s][3], e1.z, e1.w);
        }
        __syncwarp();  // every lane has consumed the stage (its values fed the mma above): safe to refill it
        if (lane == 0 && c + K3A_STAGES < c_end) issue(c + K3A_STAGES);
        if (g < H) {  // C fragment rows 0..7 = heads; columns 2t, 2t+1 = keys j0+2t, j0+2t+1
            const int ka = j0 + 2 * t, kb = j0 + 8 + 2 * t;
            if (ka < m) srow[ka] = c0[0] + d0[0];
            if (ka + 1 < m) srow[ka + 1] = c0[1] + d0[1];
            if (kb < m) srow[kb] = c1[0] + d1[0];
            if (kb + 1 < m) srow[kb + 1] = c1[1] + d1[1];
        }
    }
}

__device__ __forceinline__ float to_f32(float x) { return x; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
__device__ __forceinline__ float to_f32(__half x) { return __half2float(x); }

// Generic fp32-arithmetic variant (any C, E in fp32 or bf16): one warp per (slab row, key), lanes stride channels.
template <typename ET>
__global__ void __launch_bounds__(256)
rpe_scores_generic_kernel(const float* __restrict__ qt, long long qsb, long long qsh, long long qsn, const ET* __restrict__ E, int H, int n, int m, int C,
                          int row_begin, int row_count, long long pairs_total, float* __restrict__ scores) {
    const int lane = threadIdx.x & 31;
    const long long p = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (p >= pairs_total) return;
    const long long r = p / m;
    const int j = (int)(p - r * m);
    const int b = (int)(r / row_count);
    const int i = row_begin + (int)(r - (long long)b * row_count);
    const ET* e = E + p * C;
    for (int h = 0; h < H; ++h) {
        const float* q = qt + (size_t)b * qsb + (size_t)h * qsh + (size_t)i * qsn;
        float acc = 0.f;
        for (int c = lane; c < C; c += 32) acc = fmaf(q[c], to_f32(e[c]), acc);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (lane == 0) scores[(((size_t)b * H + h) * n + i) * m + j] = acc;
    }
}

}  // namespace geotx

using namespace geotx;

extern "C" int geotx_rpe_scores(const float* qt, int64_t qt_stride_b, int64_t qt_stride_h, int64_t qt_stride_n, const void* E, int e_dtype, int B, int H, int n, int m, int C,
                                int row_begin, int row_count, float* scores_p, void* stream) {
    GEOTX_REQUIRE(qt && E && scores_p, "geotx_rpe_scores: null pointer");
    GEOTX_REQUIRE(B > 0 && H > 0 && n > 0 && m > 0 && C > 0, "geotx_rpe_scores: bad sizes");
    GEOTX_REQUIRE(row_begin >= 0 && row_count > 0 && row_begin + row_count <= n, "geotx_rpe_scores: bad row slab");
    GEOTX_REQUIRE(e_dtype == GEOTX_F32 || e_dtype == GEOTX_BF16 || e_dtype == GEOTX_F16, "geotx_rpe_scores: bad e_dtype");
    cudaStream_t st = (cudaStream_t)stream;
    const long long rows = (long long)B * row_count;
    const bool fast = e_dtype != GEOTX_F32 && (C == 128 || C == 256) && H <= 8 &&
                      (((uintptr_t)E | (uintptr_t)qt) & 15) == 0 && qt_stride_b % 4 == 0 && qt_stride_h % 4 == 0 && qt_stride_n % 4 == 0;
    if (fast) {
        // the opt-in to > 48 KB of dynamic shared memory belongs to the device's context: once per device, not per process
        static bool opted_in[64] = {};
        int dev = 0;
        int rc = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
        if (rc) return rc;
        if (dev < 0 || dev >= 64 || !opted_in[dev]) {
            const void* kernels[4] = {(const void*)rpe_scores_tc_kernel<256, false>, (const void*)rpe_scores_tc_kernel<256, true>,
                                      (const void*)rpe_scores_tc_kernel<128, false>, (const void*)rpe_scores_tc_kernel<128, true>};
            for (int i = 0; i < 4 && !rc; ++i)
                rc = check_cuda(cudaFuncSetAttribute(kernels[i], cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                     i < 2 ? K3ACfg<256>::SMEM_BYTES : K3ACfg<128>::SMEM_BYTES),
                                "cudaFuncSetAttribute(rpe_scores_tc_kernel)");
            if (rc) return rc;
            if (dev >= 0 && dev < 64) opted_in[dev] = true;
        }
        const int num_sms = sm_count();
        // persistent: one CTA per SM, every warp streams a contiguous range of 16-key chunks
        const int chunks_per_row = (m + K3A_KEYS - 1) / K3A_KEYS;
        const long long chunks_total = rows * chunks_per_row;
        long long ctas = (chunks_total + K3A_WARPS - 1) / K3A_WARPS;
        if (ctas > num_sms) ctas = num_sms;
        const long long warps = ctas * K3A_WARPS;
        const long long chunks_per_warp = (chunks_total + warps - 1) / warps;
        const bool f16 = e_dtype == GEOTX_F16;
#define GEOTX_K3A_LAUNCH(CC, FF)                                                                                      \
    rpe_scores_tc_kernel<CC, FF><<<(unsigned)ctas, K3A_WARPS * 32, K3ACfg<CC>::SMEM_BYTES, st>>>(                        \
        qt, qt_stride_b, qt_stride_h, qt_stride_n, E, H, n, m, row_begin, row_count, chunks_total, chunks_per_row,      \
        chunks_per_warp, scores_p)
        if (C == 256) { if (f16) GEOTX_K3A_LAUNCH(256, true); else GEOTX_K3A_LAUNCH(256, false); }
        else          { if (f16) GEOTX_K3A_LAUNCH(128, true); else GEOTX_K3A_LAUNCH(128, false); }
#undef GEOTX_K3A_LAUNCH
    } else {
        long long pairs = rows * m;
        long long blocks = (pairs + 7) / 8;
        GEOTX_REQUIRE(blocks < 2147483647LL, "geotx_rpe_scores: problem too large");
        if (e_dtype == GEOTX_BF16)
            rpe_scores_generic_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(
                qt, qt_stride_b, qt_stride_h, qt_stride_n, (const __nv_bfloat16*)E, H, n, m, C, row_begin, row_count, pairs, scores_p);
        else if (e_dtype == GEOTX_F16)
            rpe_scores_generic_kernel<__half><<<(unsigned)blocks, 256, 0, st>>>(
                qt, qt_stride_b, qt_stride_h, qt_stride_n, (const __half*)E, H, n, m, C, row_begin, row_count, pairs, scores_p);
        else
            rpe_scores_generic_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(
                qt, qt_stride_b, qt_stride_h, qt_stride_n, (const float*)E, H, n, m, C, row_begin, row_count, pairs, scores_p);
    }
    return launch_status("geotx_rpe_scores");
}
