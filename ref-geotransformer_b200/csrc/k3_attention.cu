// K3b: flash-style multi-head attention with an additive pre-softmax score term.
//
// Replaces /root/reference/geotransformer/modules/transformer/rpe_transformer.py:57-70 (self blocks, scores_p from
// K3a) and /root/reference/geotransformer/modules/transformer/vanilla_transformer.py:54-68 (cross blocks,
// scores_p = NULL):   out = softmax((q k^T + scores_p) * scale  [key_mask -> -inf]) v
// The (B,H,N,M) attention matrix is never written: one CTA owns 64 queries of one (batch, head), streams 64-key
// tiles of K and V through shared memory, and keeps the running max / sum / output in registers (online
// softmax).  Both contractions run on the tensor cores as TF32 mma.sync.m16n8k8 with fp32 accumulation (error
// on a logit ~4e-4, far inside the 2e-2 budget; inputs and outputs stay fp32 in HBM).
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {

constexpr int ATT_BQ = 64;   // queries per CTA (4 warps x 16 rows)
constexpr int ATT_BK = 64;   // keys per tile
constexpr int ATT_THREADS = 128;

__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void mma_tf32_1688(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// q,k,v,out: token-major [B, tokens, H*DH]; scores_p: [B,H,n,m] or null; key_mask: [B,m] or null.
template <int DH>
__global__ void __launch_bounds__(ATT_THREADS)
attention_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                 const float* __restrict__ scores_p, const uint8_t* __restrict__ key_mask, int H, int n, int m,
                 float scale_log2e, float* __restrict__ out) {
    constexpr int PITCH = DH + 4;  // conflict-free for both fragment access patterns (see comments below)
    constexpr int KS = DH / 8;     // k-steps of q.k^T, and n-tiles of the output
    __shared__ __align__(16) float Ks[ATT_BK * PITCH];
    __shared__ __align__(16) float Vs[ATT_BK * PITCH];
    __shared__ float s_maskbias[ATT_BK];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int bh = blockIdx.y;
    const int b = bh / H, h = bh - b * H;
    const int q0 = blockIdx.x * ATT_BQ + warp * 16;
    const int ld = H * DH;
    const int r0 = q0 + g, r1 = q0 + g + 8;  // the two query rows this thread holds fragments of

    // Q fragments (A operand, row-major 16 x 8 per k-step): a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4)
    uint32_t qa[KS][4];
    {
        const float* q_r0 = q + ((size_t)b * n + min(r0, n - 1)) * ld + h * DH;
        const float* q_r1 = q + ((size_t)b * n + min(r1, n - 1)) * ld + h * DH;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
            qa[ks][0] = to_tf32(__ldg(q_r0 + 8 * ks + t));
            qa[ks][1] = to_tf32(__ldg(q_r1 + 8 * ks + t));
            qa[ks][2] = to_tf32(__ldg(q_r0 + 8 * ks + t + 4));
            qa[ks][3] = to_tf32(__ldg(q_r1 + 8 * ks + t + 4));
        }
    }
    float o[KS][4];
#pragma unroll
    for (int i = 0; i < KS; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float mx0 = -INFINITY, mx1 = -INFINITY, sum0 = 0.f, sum1 = 0.f;

    const float* sp0 = scores_p ? scores_p + (((size_t)b * H + h) * n + min(r0, n - 1)) * m : nullptr;
    const float* sp1 = scores_p ? scores_p + (((size_t)b * H + h) * n + min(r1, n - 1)) * m : nullptr;

    for (int j0 = 0; j0 < m; j0 += ATT_BK) {
        __syncthreads();  // previous tile fully consumed
        // ---- stage K and V tiles (64 x DH fp32 each), zero-filled past m ----
        for (int idx = threadIdx.x; idx < ATT_BK * (DH / 4); idx += ATT_THREADS) {
            const int row = idx / (DH / 4), c4 = idx - row * (DH / 4);
            const int j = j0 + row;
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (j < m) {
                const size_t off = ((size_t)b * m + j) * ld + h * DH + 4 * c4;
                kv = __ldg((const float4*)(k + off));
                vv = __ldg((const float4*)(v + off));
            }
            *(float4*)(Ks + row * PITCH + 4 * c4) = kv;  // PITCH % 4 == 0: 16-byte aligned
            *(float4*)(Vs + row * PITCH + 4 * c4) = vv;
        }
        if (threadIdx.x < ATT_BK) {
            const int j = j0 + threadIdx.x;
            const bool dead = j >= m || (key_mask && key_mask[(size_t)b * m + j]);
            s_maskbias[threadIdx.x] = dead ? -INFINITY : 0.f;
        }
        __syncthreads();

        // ---- S = Q K^T : 8 key n-tiles x KS k-steps.  B frag: b0 = K[key n0+g][d 8ks+t], b1 = ...[d 8ks+t+4];
        //      bank = (g*PITCH + t) mod 32 = 4g + t  -> conflict-free ----
        float s[ATT_BK / 8][4];
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            const float* kr = Ks + (nt * 8 + g) * PITCH + t;
#pragma unroll
            for (int ks = 0; ks < KS; ++ks)
                mma_tf32_1688(s[nt], qa[ks], to_tf32(kr[8 * ks]), to_tf32(kr[8 * ks + 4]));
        }
        // ---- logits (base 2), mask, running max ----
        float tmax0 = -INFINITY, tmax1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            const int jc = nt * 8 + 2 * t;  // C frag: (g, 2t) (g, 2t+1) (g+8, 2t) (g+8, 2t+1)
            float p00 = 0.f, p01 = 0.f, p10 = 0.f, p11 = 0.f;
            if (scores_p) {
                const int j = j0 + jc;
                if (j < m) { p00 = __ldg(sp0 + j); p10 = __ldg(sp1 + j); }
                if (j + 1 < m) { p01 = __ldg(sp0 + j + 1); p11 = __ldg(sp1 + j + 1); }
            }
            const float mb0 = s_maskbias[jc], mb1 = s_maskbias[jc + 1];
            s[nt][0] = (s[nt][0] + p00) * scale_log2e + mb0;
            s[nt][1] = (s[nt][1] + p01) * scale_log2e + mb1;
            s[nt][2] = (s[nt][2] + p10) * scale_log2e + mb0;
            s[nt][3] = (s[nt][3] + p11) * scale_log2e + mb1;
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
        const float corr0 = exp2f(mx0 - sub0), corr1 = exp2f(mx1 - sub1);
        mx0 = nm0; mx1 = nm1;
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            s[nt][0] = exp2f(s[nt][0] - sub0); s[nt][1] = exp2f(s[nt][1] - sub0);
            s[nt][2] = exp2f(s[nt][2] - sub1); s[nt][3] = exp2f(s[nt][3] - sub1);
            ps0 += s[nt][0] + s[nt][1];
            ps1 += s[nt][2] + s[nt][3];
        }
        sum0 = sum0 * corr0 + ps0;
        sum1 = sum1 * corr1 + ps1;
#pragma unroll
        for (int i = 0; i < KS; ++i) { o[i][0] *= corr0; o[i][1] *= corr0; o[i][2] *= corr1; o[i][3] *= corr1; }

        // ---- O += P V.  The C fragment of P (cols 2t, 2t+1) is reused directly as the A fragment by permuting
        //      the contraction index inside each 8-key group: k-slot t <-> key 2t, k-slot t+4 <-> key 2t+1.
        //      B frag: b0 = V[key 8kk+2t][d 8nt+g], b1 = V[key 8kk+2t+1][d 8nt+g];
        //      bank = (2t*PITCH + g) mod 32 = 8t + g  -> conflict-free ----
#pragma unroll
        for (int kk = 0; kk < ATT_BK / 8; ++kk) {
            uint32_t pa[4] = {to_tf32(s[kk][0]), to_tf32(s[kk][2]), to_tf32(s[kk][1]), to_tf32(s[kk][3])};
            const float* vr = Vs + (kk * 8 + 2 * t) * PITCH + g;
#pragma unroll
            for (int nt = 0; nt < KS; ++nt)
                mma_tf32_1688(o[nt], pa, to_tf32(vr[8 * nt]), to_tf32(vr[PITCH + 8 * nt]));
        }
    }
    // ---- normalise and store: 'b h n c -> b n (h c)' ----
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1);
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    const float inv0 = 1.f / sum0, inv1 = 1.f / sum1;  // all keys masked -> 0/0 = NaN, as torch.softmax gives
#pragma unroll
    for (int nt = 0; nt < KS; ++nt) {
        const int d = h * DH + nt * 8 + 2 * t;
        if (r0 < n) *(float2*)(out + ((size_t)b * n + r0) * ld + d) = make_float2(o[nt][0] * inv0, o[nt][1] * inv0);
        if (r1 < n) *(float2*)(out + ((size_t)b * n + r1) * ld + d) = make_float2(o[nt][2] * inv1, o[nt][3] * inv1);
    }
}

}  // namespace geotx

using namespace geotx;

extern "C" int geotx_attention(const float* q, const float* k, const float* v, const float* scores_p,
                               const uint8_t* key_mask, int B, int H, int n, int m, int c, float scale, float* out,
                               void* stream) {
    GEOTX_REQUIRE(q && k && v && out, "geotx_attention: null pointer");
    GEOTX_REQUIRE(B > 0 && H > 0 && n > 0 && m > 0, "geotx_attention: bad sizes");
    GEOTX_REQUIRE(c == 16 || c == 32 || c == 64, "geotx_attention: per-head dim %d not supported (16, 32, 64)", c);
    GEOTX_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)out) & 15) == 0,
                  "geotx_attention: q/k/v/out must be 16-byte aligned");
    GEOTX_REQUIRE((long long)B * H <= 65535, "geotx_attention: B*H too large");
    dim3 grid((n + ATT_BQ - 1) / ATT_BQ, B * H);
    const float sl2 = scale * 1.4426950408889634f;
    cudaStream_t st = (cudaStream_t)stream;
    if (c == 64) attention_kernel<64><<<grid, ATT_THREADS, 0, st>>>(q, k, v, scores_p, key_mask, H, n, m, sl2, out);
    else if (c == 32) attention_kernel<32><<<grid, ATT_THREADS, 0, st>>>(q, k, v, scores_p, key_mask, H, n, m, sl2, out);
    else attention_kernel<16><<<grid, ATT_THREADS, 0, st>>>(q, k, v, scores_p, key_mask, H, n, m, sl2, out);
    return launch_status("geotx_attention");
}
