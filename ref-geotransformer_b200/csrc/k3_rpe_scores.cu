// K3a: relative-position attention scores with proj_p folded into the query.
//
// Replaces /root/reference/geotransformer/modules/transformer/rpe_transformer.py:54-56
//     p = proj_p(embed_qk)                       # (B,N,M,C) x (C,C): 2 N M C^2 flop + a second N*M*C tensor
//     scores_p = einsum('bhnc,bhnmc->bhnm', q, p)
// by the exact identity
//     scores_p[b,h,i,j] = (W_p[h]^T q[b,h,i]) . E[b,i,j,:]  (+ q.b_p, constant over j: cancels in the softmax)
// so each layer reads E exactly once and never forms p.  The kernel is HBM-bound on E (2 n m C bytes in bf16):
// every warp streams rows E[b,i,j0..j0+7,:] with 16-byte loads (8 rows x 64 contiguous bytes per instruction)
// and contracts them against the folded query of row i on the tensor cores (mma.sync m16n8k16, bf16 in / fp32
// accumulate, rows of the A fragment = heads).  The folded query is split into bf16 hi + lo parts so that its
// rounding does not add to the error budget (E itself is the only bf16-rounded operand).
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {

constexpr int K3A_WARPS = 8;

__device__ __forceinline__ void mma_bf16_16816(float* c, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
    // rows 8..15 of A are zero (a1 = a3 = 0): only H <= 8 heads are populated
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// grid.x = ceil(rows * nsplit / K3A_WARPS); warp -> (slab row r, j-split sp).
// Channel permutation: lane (g = lane/4, t = lane%4) owns channels 32 s + 8 t + [0,8) of E row j0+g for every s;
// the two k16 steps of chunk s use the 8 channels as k-slots {2t,2t+1 | 2t+8,2t+9}.  Any fixed permutation of
// the contraction index is legal as long as A (the folded query) uses the same one, which it does.
template <int C>
__global__ void __launch_bounds__(K3A_WARPS * 32)
rpe_scores_bf16_kernel(const float* __restrict__ qt, long long qsb, long long qsh, const __nv_bfloat16* __restrict__ E, int H, int n, int m,
                       int row_begin, int row_count, int rows_total, int nsplit, int jchunk,
                       float* __restrict__ scores) {
    constexpr int S = C / 32;
    const int lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int wid = blockIdx.x * K3A_WARPS + (threadIdx.x >> 5);
    const int r = wid / nsplit;
    if (r >= rows_total) return;
    const int sp = wid - r * nsplit;
    const int b = r / row_count;
    const int il = r - b * row_count;   // row inside the slab
    const int i = row_begin + il;       // query row in the cloud

    // A fragments: folded query of (b, h = g, i), hi/lo bf16 split
    uint32_t ah[S][4], al[S][4];
    if (g < H) {
        const float* qrow = qt + (size_t)b * qsb + (size_t)g * qsh + (size_t)i * C + 8 * t;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const float4 u = __ldg((const float4*)(qrow + 32 * s));
            const float4 v = __ldg((const float4*)(qrow + 32 * s + 4));
            const float f[8] = {u.x, u.y, u.z, u.w, v.x, v.y, v.z, v.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float h0 = __bfloat162float(__float2bfloat16_rn(f[2 * e]));
                const float h1 = __bfloat162float(__float2bfloat16_rn(f[2 * e + 1]));
                ah[s][e] = pack_bf16x2(h0, h1);
                al[s][e] = pack_bf16x2(f[2 * e] - h0, f[2 * e + 1] - h1);
            }
        }
    } else {
#pragma unroll
        for (int s = 0; s < S; ++s)
#pragma unroll
            for (int e = 0; e < 4; ++e) { ah[s][e] = 0u; al[s][e] = 0u; }
    }

    const __nv_bfloat16* Erow = E + ((size_t)r * m) * C + 8 * t;  // E slab is [B*row_count][m][C]
    float* srow = scores + (((size_t)b * H + g) * n + i) * m;
    const int j_begin = sp * jchunk;
    const int j_end = min(m, j_begin + jchunk);

    for (int j0 = j_begin; j0 < j_end; j0 += 16) {
        uint4 e0[S], e1[S];
        const int ja = j0 + g, jb = j0 + 8 + g;
        const bool va = ja < j_end, vb = jb < j_end;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            e0[s] = va ? ldg_stream16(Erow + (size_t)ja * C + 32 * s) : make_uint4(0, 0, 0, 0);
            e1[s] = vb ? ldg_stream16(Erow + (size_t)jb * C + 32 * s) : make_uint4(0, 0, 0, 0);
        }
        float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int s = 0; s < S; ++s) {
            mma_bf16_16816(c0, ah[s][0], ah[s][1], e0[s].x, e0[s].y);
            mma_bf16_16816(c0, al[s][0], al[s][1], e0[s].x, e0[s].y);
            mma_bf16_16816(c0, ah[s][2], ah[s][3], e0[s].z, e0[s].w);
            mma_bf16_16816(c0, al[s][2], al[s][3], e0[s].z, e0[s].w);
            mma_bf16_16816(c1, ah[s][0], ah[s][1], e1[s].x, e1[s].y);
            mma_bf16_16816(c1, al[s][0], al[s][1], e1[s].x, e1[s].y);
            mma_bf16_16816(c1, ah[s][2], ah[s][3], e1[s].z, e1[s].w);
            mma_bf16_16816(c1, al[s][2], al[s][3], e1[s].z, e1[s].w);
        }
        if (g < H) {  // C fragment rows 0..7 = heads; columns 2t, 2t+1 = keys j0+2t, j0+2t+1
            const int ka = j0 + 2 * t, kb = j0 + 8 + 2 * t;
            if (ka < j_end) srow[ka] = c0[0];
            if (ka + 1 < j_end) srow[ka + 1] = c0[1];
            if (kb < j_end) srow[kb] = c1[0];
            if (kb + 1 < j_end) srow[kb + 1] = c1[1];
        }
    }
}

// Generic fp32-arithmetic variant (any C, E in fp32 or bf16): one warp per (slab row, key), lanes stride channels.
template <typename ET>
__global__ void __launch_bounds__(256)
rpe_scores_generic_kernel(const float* __restrict__ qt, long long qsb, long long qsh, const ET* __restrict__ E, int H, int n, int m, int C,
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
        const float* q = qt + (size_t)b * qsb + (size_t)h * qsh + (size_t)i * C;
        float acc = 0.f;
        for (int c = lane; c < C; c += 32) acc = fmaf(q[c], (float)e[c], acc);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (lane == 0) scores[(((size_t)b * H + h) * n + i) * m + j] = acc;
    }
}

}  // namespace geotx

using namespace geotx;

extern "C" int geotx_rpe_scores(const float* qt, int64_t qt_stride_b, int64_t qt_stride_h, const void* E, int e_dtype, int B, int H, int n, int m, int C,
                                int row_begin, int row_count, float* scores_p, void* stream) {
    GEOTX_REQUIRE(qt && E && scores_p, "geotx_rpe_scores: null pointer");
    GEOTX_REQUIRE(B > 0 && H > 0 && n > 0 && m > 0 && C > 0, "geotx_rpe_scores: bad sizes");
    GEOTX_REQUIRE(row_begin >= 0 && row_count > 0 && row_begin + row_count <= n, "geotx_rpe_scores: bad row slab");
    GEOTX_REQUIRE(e_dtype == GEOTX_F32 || e_dtype == GEOTX_BF16, "geotx_rpe_scores: bad e_dtype");
    cudaStream_t st = (cudaStream_t)stream;
    const long long rows = (long long)B * row_count;
    const bool fast = e_dtype == GEOTX_BF16 && (C == 128 || C == 256) && H <= 8 &&
                      (((uintptr_t)E | (uintptr_t)qt) & 15) == 0 && qt_stride_b % 4 == 0 && qt_stride_h % 4 == 0;
    if (fast) {
        // enough warps to keep ~32 resident per SM even for short clouds; each warp gets a contiguous key range
        int nsplit = (int)((148LL * 64 + rows - 1) / rows);
        int max_split = (m + 63) / 64;
        if (nsplit > max_split) nsplit = max_split;
        if (nsplit < 1) nsplit = 1;
        int jchunk = (((m + nsplit - 1) / nsplit) + 15) / 16 * 16;
        nsplit = (m + jchunk - 1) / jchunk;
        long long warps = rows * nsplit;
        long long blocks = (warps + K3A_WARPS - 1) / K3A_WARPS;
        GEOTX_REQUIRE(blocks < 2147483647LL && rows < 2147483647LL, "geotx_rpe_scores: problem too large");
        if (C == 256)
            rpe_scores_bf16_kernel<256><<<(unsigned)blocks, K3A_WARPS * 32, 0, st>>>(
                qt, qt_stride_b, qt_stride_h, (const __nv_bfloat16*)E, H, n, m, row_begin, row_count, (int)rows, nsplit, jchunk, scores_p);
        else
            rpe_scores_bf16_kernel<128><<<(unsigned)blocks, K3A_WARPS * 32, 0, st>>>(
                qt, qt_stride_b, qt_stride_h, (const __nv_bfloat16*)E, H, n, m, row_begin, row_count, (int)rows, nsplit, jchunk, scores_p);
    } else {
        long long pairs = rows * m;
        long long blocks = (pairs + 7) / 8;
        GEOTX_REQUIRE(blocks < 2147483647LL, "geotx_rpe_scores: problem too large");
        if (e_dtype == GEOTX_BF16)
            rpe_scores_generic_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(
                qt, qt_stride_b, qt_stride_h, (const __nv_bfloat16*)E, H, n, m, C, row_begin, row_count, pairs, scores_p);
        else
            rpe_scores_generic_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(
                qt, qt_stride_b, qt_stride_h, (const float*)E, H, n, m, C, row_begin, row_count, pairs, scores_p);
    }
    return launch_status("geotx_rpe_scores");
}
