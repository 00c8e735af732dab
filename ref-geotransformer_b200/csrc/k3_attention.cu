// K3b: flash-style multi-head attention with an additive pre-softmax score term.
//
// Replaces /root/reference/geotransformer/modules/transformer/rpe_transformer.py:57-70 (self blocks, scores_p from
// K3a) and /root/reference/geotransformer/modules/transformer/vanilla_transformer.py:54-68 (cross blocks,
// scores_p = NULL):   out = softmax((q k^T + scores_p) * scale  [key_mask -> -inf]) v
// The (B,H,N,M) attention matrix is never written: one CTA owns 64 queries of one (batch, head) and a contiguous range
// of 64-key tiles, streams K and V through a cp.async double-buffered shared-memory ring, and keeps the running max /
// sum / output in registers (online softmax).  Both contractions run on the tensor cores as TF32 mma.sync.m16n8k8 with
// fp32 accumulation (error on a logit ~4e-4, far inside the 2e-2 budget; inputs and outputs stay fp32 in HBM).
// The work per launch is small (4 n m C flop), so parallelism, not throughput, is the issue: at n = 1024 there are only
// 64 (query tile, head) pairs for 148 SMs.  The key range is therefore split over gridDim.z CTAs (flash-decoding
// style) that write unnormalised partial outputs + (max, sum) to a caller-provided workspace, merged by a second tiny
// kernel.  With one split the result is written directly.
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {

constexpr int ATT_BQ = 64;   // queries per CTA (4 warps x 16 rows)
constexpr int ATT_BK = 64;   // keys per tile
constexpr int ATT_THREADS = 128;
constexpr int ATT_TARGET_CTAS = 296;  // 2 per SM on a 148-SM B200

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
// 16-byte async copy global -> shared; src_bytes = 0 zero-fills (rows past the end of the cloud)
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int DH>
struct AttSmem {
    static constexpr int PITCH = DH + 4;  // conflict-free for both fragment access patterns (see comments below)
    static constexpr int TILE_FLOATS = ATT_BK * PITCH;
    static constexpr int BYTES = (4 * TILE_FLOATS + 2 * ATT_BK) * 4;  // K[2], V[2], maskbias[2]
};

// q,k,v,out: token-major [B, tokens, H*DH]; scores_p: [B,H,n,m] or null; key_mask: [B,m] or null.
// ws (splits > 1): float part_o[z][B*H][n][DH], then float part_ml[z][B*H][n][2] = (running max in log2 units, sum).
template <int DH>
__global__ void __launch_bounds__(ATT_THREADS)
attention_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                 const float* __restrict__ scores_p, const uint8_t* __restrict__ key_mask, int H, int n, int m, int ldq, int ldk, int ldv,
                 float scale_log2e, float* __restrict__ out, float* __restrict__ ws, int tiles_per_split) {
    using SM = AttSmem<DH>;
    constexpr int PITCH = SM::PITCH;
    constexpr int KS = DH / 8;     // k-steps of q.k^T, and n-tiles of the output
    extern __shared__ __align__(16) float att_smem[];
    float* Ks = att_smem;                            // [2][ATT_BK * PITCH]
    float* Vs = att_smem + 2 * SM::TILE_FLOATS;      // [2][ATT_BK * PITCH]
    float* s_maskbias = att_smem + 4 * SM::TILE_FLOATS;  // [2][ATT_BK]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int bh = blockIdx.y;
    const int b = bh / H, h = bh - b * H;
    const int q0 = blockIdx.x * ATT_BQ + warp * 16;
    const int ld = H * DH;
    const int r0 = q0 + g, r1 = q0 + g + 8;  // the two query rows this thread holds fragments of
    const int num_tiles = (m + ATT_BK - 1) / ATT_BK;
    const int tile_begin = blockIdx.z * tiles_per_split;
    const int tile_end = min(num_tiles, tile_begin + tiles_per_split);

    auto issue_tile = [&](int tile, int buf) {
        const int j0 = tile * ATT_BK;
        const uint32_t ks_u32 = smem_u32(Ks + buf * SM::TILE_FLOATS), vs_u32 = smem_u32(Vs + buf * SM::TILE_FLOATS);
        for (int idx = threadIdx.x; idx < ATT_BK * (DH / 4); idx += ATT_THREADS) {
            const int row = idx / (DH / 4), c4 = idx - row * (DH / 4);
            const int j = j0 + row;
            const size_t tok = (size_t)b * m + min(j, m - 1);
            const uint32_t nbytes = j < m ? 16u : 0u;
            cp_async16(ks_u32 + (row * PITCH + 4 * c4) * 4, k + tok * ldk + h * DH + 4 * c4, nbytes);  // PITCH % 4 == 0
            cp_async16(vs_u32 + (row * PITCH + 4 * c4) * 4, v + tok * ldv + h * DH + 4 * c4, nbytes);
        }
        if (threadIdx.x < ATT_BK) {
            const int j = j0 + threadIdx.x;
            const bool dead = j >= m || (key_mask && key_mask[(size_t)b * m + j]);
            s_maskbias[buf * ATT_BK + threadIdx.x] = dead ? -INFINITY : 0.f;
        }
    };

    if (tile_begin < tile_end) issue_tile(tile_begin, 0);
    cp_async_commit();

    // Q fragments (A operand, row-major 16 x 8 per k-step): a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4)
    uint32_t qa[KS][4];
    {
        const float* q_r0 = q + ((size_t)b * n + min(r0, n - 1)) * ldq + h * DH;
        const float* q_r1 = q + ((size_t)b * n + min(r1, n - 1)) * ldq + h * DH;
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
    const bool sp_vec = (m & 1) == 0;  // float2 loads need 8-byte aligned rows

    for (int tile = tile_begin; tile < tile_end; ++tile) {
        const int buf = (tile - tile_begin) & 1;
        const int j0 = tile * ATT_BK;
        if (tile + 1 < tile_end) issue_tile(tile + 1, buf ^ 1);  // that buffer was released by the trailing barrier
        cp_async_commit();
        // additive scores of this tile: issued before the wait so their latency overlaps the K/V copy
        float2 pa[ATT_BK / 8], pb[ATT_BK / 8];
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            pa[nt] = make_float2(0.f, 0.f);
            pb[nt] = make_float2(0.f, 0.f);
            if (scores_p) {
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
        __syncthreads();
        const float* Kt = Ks + buf * SM::TILE_FLOATS;
        const float* Vt = Vs + buf * SM::TILE_FLOATS;
        const float* mb = s_maskbias + buf * ATT_BK;

        // ---- S = Q K^T : 8 key n-tiles x KS k-steps.  B frag: b0 = K[key n0+g][d 8ks+t], b1 = ...[d 8ks+t+4];
        //      bank = (g*PITCH + t) mod 32 = 4g + t  -> conflict-free ----
        float s[ATT_BK / 8][4];
#pragma unroll
        for (int nt = 0; nt < ATT_BK / 8; ++nt) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            const float* kr = Kt + (nt * 8 + g) * PITCH + t;
#pragma unroll
            for (int ks = 0; ks < KS; ++ks)
                mma_tf32_1688(s[nt], qa[ks], to_tf32(kr[8 * ks]), to_tf32(kr[8 * ks + 4]));
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
            uint32_t pfrag[4] = {to_tf32(s[kk][0]), to_tf32(s[kk][2]), to_tf32(s[kk][1]), to_tf32(s[kk][3])};
            const float* vr = Vt + (kk * 8 + 2 * t) * PITCH + g;
#pragma unroll
            for (int nt = 0; nt < KS; ++nt)
                mma_tf32_1688(o[nt], pfrag, to_tf32(vr[8 * nt]), to_tf32(vr[PITCH + 8 * nt]));
        }
        __syncthreads();  // this tile's buffer may be refilled by the next iteration's issue_tile
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
        for (int nt = 0; nt < KS; ++nt) {
            const int d = h * DH + nt * 8 + 2 * t;
            if (r0 < n) *(float2*)(out + ((size_t)b * n + r0) * ld + d) = make_float2(o[nt][0] * inv0, o[nt][1] * inv0);
            if (r1 < n) *(float2*)(out + ((size_t)b * n + r1) * ld + d) = make_float2(o[nt][2] * inv1, o[nt][3] * inv1);
        }
    } else {
        const size_t rows_total = (size_t)gridDim.y * n;
        float* po = ws + ((size_t)blockIdx.z * rows_total + (size_t)bh * n) * DH;
        float* pml = ws + (size_t)gridDim.z * rows_total * DH + ((size_t)blockIdx.z * rows_total + (size_t)bh * n) * 2;
#pragma unroll
        for (int nt = 0; nt < KS; ++nt) {
            const int d = nt * 8 + 2 * t;
            if (r0 < n) *(float2*)(po + (size_t)r0 * DH + d) = make_float2(o[nt][0], o[nt][1]);
            if (r1 < n) *(float2*)(po + (size_t)r1 * DH + d) = make_float2(o[nt][2], o[nt][3]);
        }
        if (t == 0) {
            if (r0 < n) *(float2*)(pml + (size_t)r0 * 2) = make_float2(mx0, sum0);
            if (r1 < n) *(float2*)(pml + (size_t)r1 * 2) = make_float2(mx1, sum1);
        }
    }
}

// Merge of the key-range splits: out = sum_z 2^(m_z - M) O_z / sum_z 2^(m_z - M) l_z.   One thread per (row, channel).
template <int DH>
__global__ void __launch_bounds__(256)
attention_combine_kernel(const float* __restrict__ ws, int splits, int H, int n, long long rows_total,
                         float* __restrict__ out) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows_total * DH) return;
    const long long row = idx / DH;  // = (b*H + h)*n + i
    const int d = (int)(idx - row * DH);
    const float* pml = ws + (size_t)splits * rows_total * DH;
    float M = -INFINITY;
    for (int z = 0; z < splits; ++z) M = fmaxf(M, pml[((size_t)z * rows_total + row) * 2]);
    float num = 0.f, den = 0.f;
    for (int z = 0; z < splits; ++z) {
        const float mz = pml[((size_t)z * rows_total + row) * 2];
        const float w = exp2f(mz - M);  // every split masked: -inf - -inf = NaN, as torch.softmax gives
        num += w * ws[((size_t)z * rows_total + row) * DH + d];
        den += w * pml[((size_t)z * rows_total + row) * 2 + 1];
    }
    const long long bh = row / n;
    const int i = (int)(row - bh * n);
    const long long b = bh / H;
    const int h = (int)(bh - b * H);
    out[((size_t)b * n + i) * (size_t)(H * DH) + h * DH + d] = num / den;
}

}  // namespace geotx

using namespace geotx;

// Key-range splits: enough CTAs for ~2 per SM, never more than one per 64-key tile.
static void attention_plan(int B, int H, int n, int m, int* splits, int* tiles_per_split) {
    const long long base = (long long)((n + ATT_BQ - 1) / ATT_BQ) * B * H;
    const int tiles = (m + ATT_BK - 1) / ATT_BK;
    long long want = (ATT_TARGET_CTAS + base - 1) / base;
    if (want < 1) want = 1;
    if (want > tiles) want = tiles;
    *tiles_per_split = (int)((tiles + want - 1) / want);
    *splits = (tiles + *tiles_per_split - 1) / *tiles_per_split;
}

extern "C" int64_t geotx_attention_workspace_bytes(int B, int H, int n, int m, int c) {
    if (B <= 0 || H <= 0 || n <= 0 || m <= 0 || c <= 0) return 0;
    int splits, tps;
    attention_plan(B, H, n, m, &splits, &tps);
    if (splits <= 1) return 0;
    return (int64_t)splits * B * H * n * (c + 2) * 4;
}

template <int DH>
static int launch_attention(const float* q, const float* k, const float* v, const float* scores_p,
                            const uint8_t* key_mask, int B, int H, int n, int m, int ldq, int ldk, int ldv, float sl2,
                            float* out, float* ws, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        int rc = check_cuda(cudaFuncSetAttribute(attention_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 AttSmem<DH>::BYTES), "cudaFuncSetAttribute(attention_kernel)");
        if (rc) return rc;
        configured = true;
    }
    int splits = 1, tps = (m + ATT_BK - 1) / ATT_BK;
    if (ws) attention_plan(B, H, n, m, &splits, &tps);
    dim3 grid((n + ATT_BQ - 1) / ATT_BQ, B * H, splits);
    attention_kernel<DH><<<grid, ATT_THREADS, AttSmem<DH>::BYTES, st>>>(q, k, v, scores_p, key_mask, H, n, m, ldq, ldk,
                                                                         ldv, sl2, out, ws, tps);
    if (splits > 1) {
        const long long rows = (long long)B * H * n;
        const long long blocks = (rows * DH + 255) / 256;
        attention_combine_kernel<DH><<<(unsigned)blocks, 256, 0, st>>>(ws, splits, H, n, rows, out);
    }
    return launch_status("geotx_attention");
}

extern "C" int geotx_attention(const float* q, const float* k, const float* v, const float* scores_p,
                               const uint8_t* key_mask, int B, int H, int n, int m, int c, int ld_q, int ld_k,
                               int ld_v, float scale, float* out, void* workspace, void* stream) {
    GEOTX_REQUIRE(q && k && v && out, "geotx_attention: null pointer");
    GEOTX_REQUIRE(B > 0 && H > 0 && n > 0 && m > 0, "geotx_attention: bad sizes");
    GEOTX_REQUIRE(c == 16 || c == 32 || c == 64, "geotx_attention: per-head dim %d not supported (16, 32, 64)", c);
    GEOTX_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)out | (uintptr_t)workspace) & 15) == 0,
                  "geotx_attention: q/k/v/out/workspace must be 16-byte aligned");
    GEOTX_REQUIRE(scores_p == nullptr || ((uintptr_t)scores_p & 7) == 0, "geotx_attention: scores_p must be 8-byte aligned");
    GEOTX_REQUIRE((long long)B * H <= 65535, "geotx_attention: B*H too large");
    GEOTX_REQUIRE(ld_q >= H * c && ld_k >= H * c && ld_v >= H * c && ld_q % 4 == 0 && ld_k % 4 == 0 && ld_v % 4 == 0,
                  "geotx_attention: row strides must be >= H*c and multiples of 4 floats");
    const float sl2 = scale * 1.4426950408889634f;
    cudaStream_t st = (cudaStream_t)stream;
    float* ws = (float*)workspace;
    if (c == 64) return launch_attention<64>(q, k, v, scores_p, key_mask, B, H, n, m, ld_q, ld_k, ld_v, sl2, out, ws, st);
    if (c == 32) return launch_attention<32>(q, k, v, scores_p, key_mask, B, H, n, m, ld_q, ld_k, ld_v, sl2, out, ws, st);
    return launch_attention<16>(q, k, v, scores_p, key_mask, B, H, n, m, ld_q, ld_k, ld_v, sl2, out, ws, st);
}
