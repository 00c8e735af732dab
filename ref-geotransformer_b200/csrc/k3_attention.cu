// K3b: flash-style multi-head attention with an additive pre-softmax score term, for fp32 q, k, v.
// (The fused schedule of modules.py hands q, k, v over as IEEE half and runs K3h, k3_attention_f16.cu; this kernel
// serves direct callers with fp32 operands and GEOTX_ATT=f32.  Same arithmetic: it rounds q, k, v to half itself.)
//
// Replaces /root/reference/geotransformer/modules/transformer/rpe_transformer.py:57-70 (self blocks, scores_p from
// K3a) and /root/reference/geotransformer/modules/transformer/vanilla_transformer.py:54-68 (cross blocks,
// scores_p = NULL):   out = softmax((q k^T + scores_p) * scale  [key_mask -> -inf]) v
// The (B,H,N,M) attention matrix is never written: one CTA owns 64 queries of one (batch, head) and a contiguous range
// of 64-key tiles and keeps the running max / sum / output in registers (online softmax).
//
// Data path.  q, k, v arrive as fp32 (the outputs of the projection GEMMs).  K and V tiles are brought in with cp.async
// (double buffered, raw fp32), converted once per tile to IEEE half in a second shared-memory buffer, and both
// contractions run as mma.sync.m16n8k16 f16 with fp32 accumulation, operands fetched with ldmatrix (.trans for V).
// Half keeps an 11-bit significand -- the same as the TF32 the first version of this kernel used -- but needs half
// the MMAs and 1/8 of the shared-memory load instructions (the TF32 version was bound by LDS.32 + cvt, not by math).
// Logits, softmax and the output accumulator are fp32; the additive scores are added in fp32.
//
// Parallelism.  The work per launch is small (4 n m C flop): at n = 1024 there are only 64 (query tile, head) pairs
// for 148 SMs.  The key range is therefore split (flash-decoding style) over the 2, 4 or 8 CTAs of a thread-block
// CLUSTER (cluster dims (1,1,splits)): every CTA leaves its unnormalised partial output + (max, sum) in its own shared
// memory, and after one cluster barrier each CTA merges a slice of the 64 query rows by reading its peers' partials
// through distributed shared memory -- no workspace, no atomics, no second launch.  With one split the result is
// written directly.
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "sm100_ptx.cuh"
#include "host_util.h"

namespace geotx {

// Two CTA shapes (template parameter WARPS, 16 query rows per warp): 4 warps / 64 queries, two CTAs per SM -- more CTAs
// for small clouds -- and 8 warps / 128 queries, one CTA per SM -- every K/V tile is fetched from L2 and converted once
// per CTA, so the larger shape halves that traffic; it is ~8 % faster from n = 2048 up (measured, tools/k3b_bench.py).
constexpr int ATT_BK = 64;   // keys per tile
__host__ __device__ constexpr int att_ctas_per_sm(int warps) { return warps <= 4 ? 2 : 1; }  // ~210 registers / thread
constexpr int ATT_MAX_SPLITS = 8;      // portable cluster size

__device__ __forceinline__ void mma_f16_16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t addr, uint32_t& r0, uint32_t& r1) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                  uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
// two fp32 -> packed half2, saturating to the largest finite half instead of inf (inf * 0 would poison a row)
// (one F2FP.SATFINITE.F16.F32.PACK_AB)
__device__ __forceinline__ uint32_t pack_half2_sat(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
// 2^x, one MUFU.EX2 (exp2f() adds range fix-ups that the softmax does not need: arguments are <= 0, -inf -> 0)
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// 16-byte async copy global -> shared; src_bytes = 0 zero-fills (rows past the end of the cloud)
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int DH>
struct AttSmem {
    static constexpr int STAGE_FLOATS = ATT_BK * DH;  // one raw fp32 tile (K or V), rows dense
    static constexpr int HPITCH = DH + 8;             // halves per row of the converted tiles: 16-byte aligned rows whose
                                                      // pitch / 16 B is odd -> ldmatrix is bank-conflict free
    static constexpr int HTILE_HALVES = ATT_BK * HPITCH;
    // raw K[2], raw V[2], half K, half V, maskbias[2]
    static constexpr int BYTES = 4 * STAGE_FLOATS * 4 + 2 * HTILE_HALVES * 2 + 2 * ATT_BK * 4;
    // key-split merge: the raw tiles are dead by then; partial O [64][PO_PITCH] + (max, sum)[64] live in their place
    static constexpr int PO_PITCH = DH + 4;
    static_assert((128 * PO_PITCH + 128 * 2) * 4 <= 4 * STAGE_FLOATS * 4, "partials must fit in the raw tiles");
};

struct AttParams {
    const float* q;
    const float* k;
    const float* v;
    const float* scores_p;
    const uint8_t* key_mask;
    float* out;
    int H, n, m, ldq, ldk, ldv;
    float scale_log2e;
    int tiles_per_split;
};

template <int DH, int WARPS>
__global__ void __launch_bounds__(32 * WARPS, att_ctas_per_sm(WARPS))
attention_kernel(const __grid_constant__ AttParams prm) {
    constexpr int ATT_BQ = 16 * WARPS, ATT_THREADS = 32 * WARPS;
    using SM = AttSmem<DH>;
    constexpr int HP = SM::HPITCH;
    constexpr int KS = DH / 16;  // k16 steps of q.k^T
    constexpr int NT = DH / 8;   // 8-wide n-tiles of the output
    extern __shared__ __align__(16) uint8_t att_smem[];
    float* rawK = (float*)att_smem;               // [2][ATT_BK * DH]
    float* rawV = rawK + 2 * SM::STAGE_FLOATS;    // [2][ATT_BK * DH]
    __half* Kh = (__half*)(rawV + 2 * SM::STAGE_FLOATS);  // [ATT_BK][HP]
    __half* Vh = Kh + SM::HTILE_HALVES;                    // [ATT_BK][HP]
    float* s_maskbias = (float*)(Vh + SM::HTILE_HALVES);   // [2][ATT_BK]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int H = prm.H, n = prm.n, m = prm.m;
    const int bh = blockIdx.y;
    const int b = bh / H, h = bh - b * H;
    const int q0 = blockIdx.x * ATT_BQ + warp * 16;
    const int ld = H * DH;
    const int r0 = q0 + g, r1 = q0 + g + 8;  // the two query rows this thread holds fragments of
    const int num_tiles = (m + ATT_BK - 1) / ATT_BK;
    const int tile_begin = blockIdx.z * prm.tiles_per_split;
    const int tile_end = min(num_tiles, tile_begin + prm.tiles_per_split);
    const uint32_t rawK_u32 = smem_u32(rawK), rawV_u32 = smem_u32(rawV);
    const uint32_t Kh_u32 = smem_u32(Kh), Vh_u32 = smem_u32(Vh);

    auto issue_tile = [&](int tile, int buf) {
        const int j0 = tile * ATT_BK;
        const uint32_t kdst = rawK_u32 + buf * SM::STAGE_FLOATS * 4, vdst = rawV_u32 + buf * SM::STAGE_FLOATS * 4;
        for (int idx = threadIdx.x; idx < ATT_BK * (DH / 4); idx += ATT_THREADS) {
            const int row = idx / (DH / 4), c4 = idx - row * (DH / 4);
            const int j = j0 + row;
            const size_t tok = (size_t)b * m + min(j, m - 1);
            const uint32_t nbytes = j < m ? 16u : 0u;
            cp_async16(kdst + idx * 16, prm.k + tok * prm.ldk + h * DH + 4 * c4, nbytes);
            cp_async16(vdst + idx * 16, prm.v + tok * prm.ldv + h * DH + 4 * c4, nbytes);
        }
        if (threadIdx.x < ATT_BK) {
            const int j = j0 + threadIdx.x;
            const bool dead = j >= m || (prm.key_mask && prm.key_mask[(size_t)b * m + j]);
            s_maskbias[buf * ATT_BK + threadIdx.x] = dead ? -INFINITY : 0.f;
        }
    };

    if (tile_begin < tile_end) issue_tile(tile_begin, 0);
    cp_async_commit();

    // Q fragments (A operand, 16 x 16 per k-step): a0 (g, 2t..) a1 (g+8, 2t..) a2 (g, 2t+8..) a3 (g+8, 2t+8..)
    uint32_t qa[KS][4];
    {
        const float* q_r0 = prm.q + ((size_t)b * n + min(r0, n - 1)) * prm.ldq + h * DH + 2 * t;
        const float* q_r1 = prm.q + ((size_t)b * n + min(r1, n - 1)) * prm.ldq + h * DH + 2 * t;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
            const float2 x0 = __ldg((const float2*)(q_r0 + 16 * ks)), x1 = __ldg((const float2*)(q_r1 + 16 * ks));
            const float2 x2 = __ldg((const float2*)(q_r0 + 16 * ks + 8)), x3 = __ldg((const float2*)(q_r1 + 16 * ks + 8));
            qa[ks][0] = pack_half2_sat(x0.x, x0.y);
            qa[ks][1] = pack_half2_sat(x1.x, x1.y);
            qa[ks][2] = pack_half2_sat(x2.x, x2.y);
            qa[ks][3] = pack_half2_sat(x3.x, x3.y);
        }
    }
    float o[NT][4];
#pragma unroll
    for (int i = 0; i < NT; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float mx0 = -INFINITY, mx1 = -INFINITY, sum0 = 0.f, sum1 = 0.f;

    const float* sp0 = prm.scores_p ? prm.scores_p + (((size_t)b * H + h) * n + min(r0, n - 1)) * m : nullptr;
    const float* sp1 = prm.scores_p ? prm.scores_p + (((size_t)b * H + h) * n + min(r1, n - 1)) * m : nullptr;
    const bool sp_vec = (m & 1) == 0;  // float2 loads need 8-byte aligned rows
    const float scale_log2e = prm.scale_log2e;

    for (int tile = tile_begin; tile < tile_end; ++tile) {
        const int buf = (tile - tile_begin) & 1;
        const int j0 = tile * ATT_BK;
        if (tile + 1 < tile_end) issue_tile(tile + 1, buf ^ 1);  // that raw buffer was drained by the last conversion
        cp_async_commit();
        // additive scores of this tile: issued before the wait so their latency overlaps the K/V copy
        float2 pa[ATT_BK / 8], pb[ATT_BK / 8];
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            pa[nt] = make_float2(0.f, 0.f);
            pb[nt] = make_float2(0.f, 0.f);
            if (sp0) {
                const int j = j0 + nt * 8 + 2 * t;  // C frag: (g, 2t) (g, 2t+1) (g+8, 2t) (g+8, 2t+1)
                if (sp_vec) {
                    if (j < m) { pa[nt] = __ldg((const float2*)(sp0 + j)); pb[nt] = __ldg((const float2*)(sp1 + j)); }
                } else {
                    if (j < m) { pa[nt].x = __ldg(sp0 + j); pb[nt].x = __ldg(sp1 + j); }
                    if (j + 1 < m) { pa[nt].y = __ldg(sp0 + j + 1); pb[nt].y = __ldg(sp1 + j + 1); }
                }
            }
        }
        cp_async_wait<1>();   // everything but the group just committed has landed
        __syncthreads();      // ... for every thread; and nobody still reads the half tiles of the previous tile
        // ---- fp32 -> half, once per tile (rows keep their order; pitch HP) ----
        {
            const float4* rk = (const float4*)(rawK + buf * SM::STAGE_FLOATS);
            const float4* rv = (const float4*)(rawV + buf * SM::STAGE_FLOATS);
#pragma unroll 4
            for (int idx = threadIdx.x; idx < ATT_BK * (DH / 4); idx += ATT_THREADS) {
                const int row = idx / (DH / 4), c4 = idx - row * (DH / 4);
                const float4 kk = rk[idx], vv = rv[idx];
                *(uint2*)(Kh + row * HP + 4 * c4) = make_uint2(pack_half2_sat(kk.x, kk.y), pack_half2_sat(kk.z, kk.w));
                *(uint2*)(Vh + row * HP + 4 * c4) = make_uint2(pack_half2_sat(vv.x, vv.y), pack_half2_sat(vv.z, vv.w));
            }
        }
        __syncthreads();
        const float* mb = s_maskbias + buf * ATT_BK;

        // ---- S = Q K^T : 8 key n-tiles x KS k-steps.  B frag (k = channel, n = key) straight from the row-major
        //      half tile: ldmatrix matrix j = channels 8j..8j+7 of keys 8nt..8nt+7 ----
        float s[ATT_BK / 8][4];
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            if (KS >= 2) {
#pragma unroll
                for (int ks2 = 0; ks2 < KS / 2; ++ks2) {
                    uint32_t b0, b1, b2, b3;
                    ldmatrix_x4(Kh_u32 + ((nt * 8 + (lane & 7)) * HP + 32 * ks2 + 8 * (lane >> 3)) * 2, b0, b1, b2, b3);
                    mma_f16_16816(s[nt], qa[2 * ks2], b0, b1);
                    mma_f16_16816(s[nt], qa[(2 * ks2 + 1) % KS], b2, b3);
                }
            } else {
                uint32_t b0, b1;
                ldmatrix_x2(Kh_u32 + ((nt * 8 + (lane & 7)) * HP + 8 * ((lane >> 3) & 1)) * 2, b0, b1);
                mma_f16_16816(s[nt], qa[0], b0, b1);
            }
        }
        // ---- logits (base 2), mask, running max ----
        float tmax0 = -INFINITY, tmax1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            const int jc = nt * 8 + 2 * t;
            const float mb0 = mb[jc], mb1 = mb[jc + 1];
            s[nt][0] = (s[nt][0] + pa[nt].x) * scale_log2e + mb0;
            s[nt][1] = (s[nt][1] + pa[nt].y) * scale_log2e + mb1;
            s[nt][2] = (s[nt][2] + pb[nt].x) * scale_log2e + mb0;
            s[nt][3] = (s[nt][3] + pb[nt].y) * scale_log2e + mb1;
            tmax0 = fmaxf(tmax0, fmaxf(s[nt][0], s[nt][1]));
            tmax1 = fmaxf(tmax1, fmaxf(s[nt][2], s[nt][3]));
        }
        tmax0 = fmaxf(tmax0, __shfl_xor_sync(0xffffffffu, tmax0, 1));
        tmax0 = fmaxf(tmax0, __shfl_xor_sync(0xffffffffu, tmax0, 2));
        tmax1 = fmaxf(tmax1, __shfl_xor_sync(0xffffffffu, tmax1, 1));
        tmax1 = fmaxf(tmax1, __shfl_xor_sync(0xffffffffu, tmax1, 2));
        const float nm0 = fmaxf(mx0, tmax0), nm1 = fmaxf(mx1, tmax1);
        // a row whose keys so far are all masked keeps max = -inf; use 0 as the subtrahend to avoid inf - inf
        const float sub0 = nm0 == -INFINITY ? 0.f : nm0, sub1 = nm1 == -INFINITY ? 0.f : nm1;
        const float corr0 = fast_exp2(mx0 - sub0), corr1 = fast_exp2(mx1 - sub1);
        mx0 = nm0; mx1 = nm1;
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            s[nt][0] = fast_exp2(s[nt][0] - sub0); s[nt][1] = fast_exp2(s[nt][1] - sub0);
            s[nt][2] = fast_exp2(s[nt][2] - sub1); s[nt][3] = fast_exp2(s[nt][3] - sub1);
            ps0 += s[nt][0] + s[nt][1];
            ps1 += s[nt][2] + s[nt][3];
        }
        sum0 = sum0 * corr0 + ps0;
        sum1 = sum1 * corr1 + ps1;
#pragma unroll
        for (int i = 0; i < NT; ++i) { o[i][0] *= corr0; o[i][1] *= corr0; o[i][2] *= corr1; o[i][3] *= corr1; }

        // ---- O += P V.  The C fragments of two adjacent key n-tiles are exactly the A fragment of one k16 step.
        //      B frag (k = key, n = channel) needs V transposed: ldmatrix.trans on the row-major half tile;
        //      matrix j: keys 16kk + 8(j&1).., channels 16nt2 + 8(j>>1).. ----
#pragma unroll
        for (int kk = 0; kk < ATT_BK / 16; ++kk) {
            uint32_t pf[4];
            {
                const __half2 h0 = __floats2half2_rn(s[2 * kk][0], s[2 * kk][1]);
                const __half2 h1 = __floats2half2_rn(s[2 * kk][2], s[2 * kk][3]);
                const __half2 h2 = __floats2half2_rn(s[2 * kk + 1][0], s[2 * kk + 1][1]);
                const __half2 h3 = __floats2half2_rn(s[2 * kk + 1][2], s[2 * kk + 1][3]);
                pf[0] = *(const uint32_t*)&h0; pf[1] = *(const uint32_t*)&h1;
                pf[2] = *(const uint32_t*)&h2; pf[3] = *(const uint32_t*)&h3;
            }
            const int key = 16 * kk + 8 * ((lane >> 3) & 1) + (lane & 7);
#pragma unroll
            for (int nt2 = 0; nt2 < NT / 2; ++nt2) {
                uint32_t b0, b1, b2, b3;
                ldmatrix_x4_trans(Vh_u32 + (key * HP + 16 * nt2 + 8 * (lane >> 4)) * 2, b0, b1, b2, b3);
                mma_f16_16816(o[2 * nt2], pf, b0, b1);
                mma_f16_16816(o[2 * nt2 + 1], pf, b2, b3);
            }
        }
    }
    cp_async_wait<0>();

    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1);
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    if (gridDim.z == 1) {
        // ---- normalise and store: 'b h n c -> b n (h c)' ----
        const float inv0 = 1.f / sum0, inv1 = 1.f / sum1;  // all keys masked -> 0/0 = NaN, as torch.softmax gives
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int d = h * DH + nt * 8 + 2 * t;
            if (r0 < n) *(float2*)(prm.out + ((size_t)b * n + r0) * ld + d) = make_float2(o[nt][0] * inv0, o[nt][1] * inv0);
            if (r1 < n) *(float2*)(prm.out + ((size_t)b * n + r1) * ld + d) = make_float2(o[nt][2] * inv1, o[nt][3] * inv1);
        }
        return;
    }

    // ---- split keys: partial O and (max, sum) stay in this CTA's shared memory; the cluster merges them ----
    constexpr int POP = SM::PO_PITCH;
    float* po = (float*)att_smem;          // [ATT_BQ][POP]   (the raw K/V tiles are no longer needed)
    float* pml = po + ATT_BQ * POP;        // [ATT_BQ][2]
    __syncthreads();                       // every warp is done with the tiles this overwrites
    {
        const int lr0 = warp * 16 + g, lr1 = lr0 + 8;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int d = nt * 8 + 2 * t;
            *(float2*)(po + lr0 * POP + d) = make_float2(o[nt][0], o[nt][1]);
            *(float2*)(po + lr1 * POP + d) = make_float2(o[nt][2], o[nt][3]);
        }
        if (t == 0) {
            *(float2*)(pml + lr0 * 2) = make_float2(mx0, sum0);
            *(float2*)(pml + lr1 * 2) = make_float2(mx1, sum1);
        }
    }
    ptx::cluster_sync();  // all partials of this (query tile, head) are in place, cluster-wide
    {
        // out = sum_z 2^(m_z - M) O_z / sum_z 2^(m_z - M) l_z.  CTA z of the cluster merges rows [z, z+1) * 64/splits;
        // one thread per (row, 4 channels), peers read through shared::cluster addresses.
        const int splits = gridDim.z;
        const int rows_per_cta = ATT_BQ / splits;
        const uint32_t po_u32 = smem_u32(po), pml_u32 = smem_u32(pml);
        for (int item = threadIdx.x; item < rows_per_cta * (DH / 4); item += ATT_THREADS) {
            const int lr = blockIdx.z * rows_per_cta + item / (DH / 4);
            const int c4 = item % (DH / 4);
            const int row = blockIdx.x * ATT_BQ + lr;
            float mz[ATT_MAX_SPLITS], lz[ATT_MAX_SPLITS];
            float M = -INFINITY;
#pragma unroll
            for (int z = 0; z < ATT_MAX_SPLITS; ++z) {
                if (z < splits) {
                    const uint32_t a = ptx::map_to_cta(pml_u32 + lr * 8, z);
                    asm volatile("ld.shared::cluster.v2.f32 {%0,%1}, [%2];" : "=f"(mz[z]), "=f"(lz[z]) : "r"(a));
                    M = fmaxf(M, mz[z]);
                }
            }
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float den = 0.f;
#pragma unroll
            for (int z = 0; z < ATT_MAX_SPLITS; ++z) {
                if (z < splits) {
                    const float w = fast_exp2(mz[z] - M);  // every split masked: -inf - -inf = NaN, as torch.softmax gives
                    den += w * lz[z];
                    const uint32_t a = ptx::map_to_cta(po_u32 + (lr * POP + 4 * c4) * 4, z);
                    float4 x;
                    asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
                                 : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "r"(a));
                    acc.x += w * x.x; acc.y += w * x.y; acc.z += w * x.z; acc.w += w * x.w;
                }
            }
            if (row < n) {
                const float inv = 1.f / den;
                *(float4*)(prm.out + ((size_t)b * n + row) * ld + h * DH + 4 * c4) =
                    make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
            }
        }
    }
    ptx::cluster_sync();  // nobody leaves while a peer may still read its shared memory
}

}  // namespace geotx

using namespace geotx;

// Key-range splits (= cluster size): the largest power of two <= 8 that keeps the grid within one wave and leaves
// every CTA at least one 64-key tile.
static void attention_plan(int B, int H, int n, int m, int warps, int allow_split, int* splits, int* tiles_per_split) {
    const int bq = 16 * warps;
    const long long base = (long long)((n + bq - 1) / bq) * B * H;
    const int tiles = (m + ATT_BK - 1) / ATT_BK;
    const int wave = sm_count() * att_ctas_per_sm(warps);
    int s = 1;
    if (allow_split)  // stay within one wave: a second, partly filled wave costs more than the split gains
        while (s < ATT_MAX_SPLITS && base * (2 * s) <= wave && 2 * s <= tiles) s *= 2;
    *tiles_per_split = (tiles + s - 1) / s;
    while (s > 1 && (s - 1) * *tiles_per_split >= tiles) {  // no CTA of the cluster may be left without keys
        s /= 2;
        *tiles_per_split = (tiles + s - 1) / s;
    }
    *splits = s;
}

template <int DH, int WARPS>
static int launch_attention(AttParams prm, int B, int allow_split, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        int rc = check_cuda(cudaFuncSetAttribute(attention_kernel<DH, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 AttSmem<DH>::BYTES), "cudaFuncSetAttribute(attention_kernel)");
        if (rc) return rc;
        configured = true;
    }
    int splits, tps;
    attention_plan(B, prm.H, prm.n, prm.m, WARPS, allow_split, &splits, &tps);
    prm.tiles_per_split = tps;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((prm.n + 16 * WARPS - 1) / (16 * WARPS), B * prm.H, splits);
    cfg.blockDim = dim3(32 * WARPS, 1, 1);
    cfg.dynamicSmemBytes = AttSmem<DH>::BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = splits;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int rc = check_cuda(cudaLaunchKernelEx(&cfg, attention_kernel<DH, WARPS>, prm), "cudaLaunchKernelEx(attention_kernel)");
    if (rc) return rc;
    return launch_status("geotx_attention");
}

template <int DH>
static int launch_attention_shape(const AttParams& prm, int B, int allow_split, cudaStream_t st) {
    // 64-query CTAs (4 warps, two per SM) up to n = 2048 at H = 4, 128-query CTAs (8 warps, one per SM) above.  Alone,
    // the large shape is 2-8 % faster from n = 1024 up (K/V tiles fetched and converted once per 128 queries), but a
    // serving job keeps several pairs in flight and then SM-time counts: with the small shape the step is 3.5 % / 2 %
    // shorter at n = 1024 / 2048 (bench.py, 4 graph instances: 0.804 -> 0.777 ms, 2.718 -> 2.666 ms)
    const long long big = (long long)((prm.n + 127) / 128) * B * prm.H;
#ifndef ATT_BIG_THRESHOLD
#define ATT_BIG_THRESHOLD 128  // (B H n / 128 query tiles) from which the 128-query CTA shape is used
#endif
    if (big >= ATT_BIG_THRESHOLD) return launch_attention<DH, 8>(prm, B, allow_split, st);
    return launch_attention<DH, 4>(prm, B, allow_split, st);
}

extern "C" int geotx_attention(const float* q, const float* k, const float* v, const float* scores_p,
                               const uint8_t* key_mask, int B, int H, int n, int m, int c, int ld_q, int ld_k,
                               int ld_v, float scale, float* out, int split_keys, void* stream) {
    GEOTX_REQUIRE(q && k && v && out, "geotx_attention: null pointer");
    GEOTX_REQUIRE(B > 0 && H > 0 && n > 0 && m > 0, "geotx_attention: bad sizes");
    GEOTX_REQUIRE(c == 16 || c == 32 || c == 64, "geotx_attention: per-head dim %d not supported (16, 32, 64)", c);
    GEOTX_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)out) & 15) == 0,
                  "geotx_attention: q/k/v/out must be 16-byte aligned");
    GEOTX_REQUIRE(scores_p == nullptr || ((uintptr_t)scores_p & 7) == 0, "geotx_attention: scores_p must be 8-byte aligned");
    GEOTX_REQUIRE((long long)B * H <= 65535, "geotx_attention: B*H too large");
    GEOTX_REQUIRE(ld_q >= H * c && ld_k >= H * c && ld_v >= H * c && ld_q % 4 == 0 && ld_k % 4 == 0 && ld_v % 4 == 0,
                  "geotx_attention: row strides must be >= H*c and multiples of 4 floats");
    AttParams prm;
    prm.q = q; prm.k = k; prm.v = v; prm.scores_p = scores_p; prm.key_mask = key_mask; prm.out = out;
    prm.H = H; prm.n = n; prm.m = m; prm.ldq = ld_q; prm.ldk = ld_k; prm.ldv = ld_v;
    prm.scale_log2e = scale * 1.4426950408889634f;
    prm.tiles_per_split = 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (c == 64) return launch_attention_shape<64>(prm, B, split_keys, st);
    if (c == 32) return launch_attention_shape<32>(prm, B, split_keys, st);
    return launch_attention_shape<16>(prm, B, split_keys, st);
}
