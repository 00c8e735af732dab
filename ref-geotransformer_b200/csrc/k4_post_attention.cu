// K4f: everything a transformer layer does after its attention core, in ONE kernel:
//     x1  = LayerNorm(linear(hidden) + x_in)                       rpe_transformer.py:101-103 / vanilla_transformer.py:99-101
//     out = LayerNorm(x1 + squeeze(ReLU(expand(x1))))              output_layer.py:15-21 (AttentionOutput)
// The reference (and round 1 of this repo) runs these as 3 library GEMMs + 2 residual/LayerNorm kernels per layer and
// cloud step: 45 launches of a few microseconds per forward, each one latency-bound at ~2k tokens.  All five steps are
// token-local (LayerNorm normalises over the channels of ONE token), so a CTA that owns a tile of 32 tokens runs the
// whole chain without leaving the SM: the activations live in shared memory between the GEMMs (half as mma A operands,
// fp32 for the residuals), the contractions run on mma.sync.m16n8k16 (f16 operands -- the same 11-bit significand as
// the TF32 library path it replaces -- fp32 accumulation), LayerNorm is fp32 with warp shuffles (torch.nn.LayerNorm:
// biased variance, eps inside the root).
//
// The weights (655 KB of halves per CTA at C = 256, from L2) arrive through the TMA unit: they are packed once per
// weight set (geotx_pack_linear_weight) into K chunks of 32 whose rows are 64 dense bytes with the 16-byte units
// XOR-swizzled for conflict-free ldmatrix, so that chunk c of a matrix is ONE contiguous block and ONE cp.async.bulk
// issued by a dedicated loader warp into a 4-stage ring; the eight compute warps wait on the stage's mbarrier
// (complete_tx), contract, and hand the stage back with one arrival per warp.  No CTA-wide barrier inside a GEMM.
// The first version issued 16-byte cp.async from the compute warps themselves and put a __syncthreads around every
// chunk: the copy-issue phase (64 LSU instructions per chunk) and the mma phase of all eight warps alternated instead of
// overlapping, 0.9 us per chunk and 32 us per launch for 5 us of tensor work (profiles/r02_launches_3dmatch_1024.csv).
// Why mma.sync and not tcgen05: at T ~ 2k tokens a 128-token tcgen05 tile gives 16 CTAs on 148 SMs, and every CTA has
// to pull all three weight matrices through its own L2 port; 32-token tiles put the same chain on 64 SMs.
#include <cuda_fp16.h>
#include <type_traits>
#include "geotx_common.cuh"
#include "host_util.h"
#include "sm100_ptx.cuh"
#include "dense_tile.cuh"

namespace geotx {

struct PostAttnArgs {
    const float* hidden; long long ld_hidden;   // attention output [T, C]
    const float* x_in; long long ld_x;          // the layer's input tokens (residual) [T, C]
    float* out; long long ld_out;               // [T, C]
    const __half* w_lin; const float* b_lin;    // packed [C][C]
    const float* g1; const float* be1;          // attention.norm
    const __half* w_exp; const float* b_exp;    // packed [2C][C]
    const __half* w_sq; const float* b_sq;      // packed [C][2C]
    const float* g2; const float* be2;          // output.norm
    float eps1, eps2;
    int T;
};

template <int C>
struct PACfg {
    static constexpr int A_PITCH = C * 2 + 16;        // half A operand [32][C]      (bytes, odd multiple of 16)
    static constexpr int E_PITCH = 2 * C * 2 + 16;    // half A operand [32][2C]
    static constexpr int X_PITCH = C + 4;             // fp32 staging / residual [32][C] (floats)
    static constexpr int STAGE_BYTES = 2 * C * DT_ROW_BYTES;  // the widest GEMM (N = 2C) sizes a stage
    static constexpr int OFF_BAR = 0;                 // full[DT_STAGES], empty[DT_STAGES]
    static constexpr int OFF_A = 128;
    static constexpr int OFF_E = OFF_A + DT_BM * A_PITCH;
    static constexpr int OFF_X = OFF_E + DT_BM * E_PITCH;
    static constexpr int OFF_W = (OFF_X + DT_BM * X_PITCH * 4 + 127) / 128 * 128;
    static constexpr int SMEM = OFF_W + DT_STAGES * STAGE_BYTES;
    // ring schedule: GEMM 0 (linear: N = C, K = C), 1 (expand: N = 2C, K = C), 2 (squeeze: N = C, K = 2C).  A ring entry
    // is one stage = 2C rows x one K chunk for GEMM 1, C rows x TWO K chunks for GEMMs 0 and 2 (same bytes, half the
    // waits / fences / releases: a chunk of an N = C GEMM cost 500 clocks against 256 of HMMA, tools/k4f_phases.py)
    static constexpr int CH0 = C / DT_KC / 2, CH1 = C / DT_KC, CH2 = 2 * C / DT_KC / 2, CHUNKS = CH0 + CH1 + CH2;
};

#ifdef PA_PHASE_CLOCKS  // timing experiment: clock64 of CTA 0 / warp 0 at the phase boundaries (tools/k4f_phases.py)
__device__ long long pa_phase_clock[16];
#define PA_MARK(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) pa_phase_clock[i] = clock64(); } while (0)
#else
#define PA_MARK(i) do { } while (0)
#endif

template <int C>
__global__ void __launch_bounds__(DT_THREADS, 1) post_attention_kernel(const PostAttnArgs p) {
    using Cfg = PACfg<C>;
    extern __shared__ __align__(128) unsigned char pa_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const int row0 = blockIdx.x * DT_BM;
    const uint32_t s_base = smem_u32(pa_smem);
    const uint32_t sA = s_base + Cfg::OFF_A, sE = s_base + Cfg::OFF_E;
    float* const sX = reinterpret_cast<float*>(pa_smem + Cfg::OFF_X);
    DtRing ring;
    ring.bar_full = s_base + Cfg::OFF_BAR;
    ring.bar_empty = ring.bar_full + 8 * DT_STAGES;
    ring.stages = s_base + Cfg::OFF_W;
    ring.stage_bytes = Cfg::STAGE_BYTES;
    PA_MARK(0);
    ring.init(tid);
    __syncthreads();

    if (warp == DT_WARPS) {  // ---- loader: the three matrices' chunks, in schedule order, one bulk copy each ----
        if (lane == 0) {
            for (int c = 0; c < Cfg::CHUNKS; ++c) {
                if (c < Cfg::CH0) ring.load(c, p.w_lin + (size_t)c * 2 * C * DT_KC, 2 * C * DT_ROW_BYTES);
                else if (c < Cfg::CH0 + Cfg::CH1) ring.load(c, p.w_exp + (size_t)(c - Cfg::CH0) * 2 * C * DT_KC, 2 * C * DT_ROW_BYTES);
                else ring.load(c, p.w_sq + (size_t)(c - Cfg::CH0 - Cfg::CH1) * 2 * C * DT_KC, 2 * C * DT_ROW_BYTES);
            }
        }
        return;
    }

    // Prologue, one L2 round trip for everything: the residual rows x_in go straight into sX (cp.async; GEMM 0's epilogue
    // adds them there), and the tile's attention output is fetched with all of a thread's loads in flight before the first
    // conversion to the half A operand (a load -> convert -> store loop pays the L2 latency once per iteration: 14 % of
    // the kernel's stall samples sat on its first F2FP).
    constexpr int C4 = C / 4, PER_THREAD = DT_BM * C4 / (DT_WARPS * 32);
    static_assert(DT_BM * C4 % (DT_WARPS * 32) == 0, "tile loads divide evenly over the compute threads");
#pragma unroll
    for (int u = 0; u < PER_THREAD; ++u) {
        const int e = tid + u * DT_WARPS * 32, r = e / C4, c4 = e - r * C4;
        const int rr = min(row0 + r, p.T - 1);  // rows past the end: any valid row (their results are never stored)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s_base + Cfg::OFF_X + (r * Cfg::X_PITCH + c4 * 4) * 4),
                     "l"(p.x_in + (size_t)rr * p.ld_x + c4 * 4) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    {
        float4 v[PER_THREAD];
#pragma unroll
        for (int u = 0; u < PER_THREAD; ++u) {
            const int e = tid + u * DT_WARPS * 32, r = e / C4, c4 = e - r * C4;
            v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (row0 + r < p.T) v[u] = __ldg(reinterpret_cast<const float4*>(p.hidden + (size_t)(row0 + r) * p.ld_hidden) + c4);
        }
#pragma unroll
        for (int u = 0; u < PER_THREAD; ++u) {
            const int e = tid + u * DT_WARPS * 32, r = e / C4, c4 = e - r * C4;
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(sA + r * Cfg::A_PITCH + c4 * 8), "r"(dt_pack(v[u].x, v[u].y)),
                         "r"(dt_pack(v[u].z, v[u].w)) : "memory");
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    dt_sync_compute();
    PA_MARK(1);

    // LayerNorm of the 32 rows staged in sX (fp32): warp w normalises rows 4w .. 4w+3, lane owns C / 32 channels.  The
    // four rows go through their two shuffle reductions TOGETHER (four independent chains per level instead of one: the
    // row-at-a-time form cost 2500 clocks per LayerNorm, 18 % of the CTA's time, tools/k4f_phases.py), and gamma / beta
    // are fetched once per LayerNorm, BEFORE the GEMM in front of it (ln_load), so that their L2 latency is not paid here.
    constexpr int PER = C / 32, ROWS = DT_BM / 8;
    struct LnW { float4 gm[PER / 4], bt[PER / 4]; };
    auto ln_load = [&](const float* gamma, const float* beta, LnW& w) {
#pragma unroll
        for (int i = 0; i < PER / 4; ++i) {
            w.gm[i] = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * i);
            w.bt[i] = __ldg(reinterpret_cast<const float4*>(beta) + lane + 32 * i);
        }
    };
    auto layer_norm = [&](const LnW& w, float eps, bool to_global) {
        const float4 (&gm)[PER / 4] = w.gm;
        const float4 (&bt)[PER / 4] = w.bt;
        float v[ROWS][PER], sum[ROWS], sq[ROWS];
#pragma unroll
        for (int rr = 0; rr < ROWS; ++rr) {
            const float* row = sX + (warp * ROWS + rr) * Cfg::X_PITCH;
            sum[rr] = 0.f;
#pragma unroll
            for (int i = 0; i < PER / 4; ++i) {
                const float4 t = *reinterpret_cast<const float4*>(row + (lane + 32 * i) * 4);
                v[rr][4 * i] = t.x; v[rr][4 * i + 1] = t.y; v[rr][4 * i + 2] = t.z; v[rr][4 * i + 3] = t.w;
                sum[rr] += (t.x + t.y) + (t.z + t.w);
            }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
#pragma unroll
            for (int rr = 0; rr < ROWS; ++rr) sum[rr] += __shfl_xor_sync(0xffffffffu, sum[rr], off);
#pragma unroll
        for (int rr = 0; rr < ROWS; ++rr) {
            sum[rr] = sum[rr] / (float)C;  // mean
            sq[rr] = 0.f;
#pragma unroll
            for (int i = 0; i < PER; ++i) sq[rr] += (v[rr][i] - sum[rr]) * (v[rr][i] - sum[rr]);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
#pragma unroll
            for (int rr = 0; rr < ROWS; ++rr) sq[rr] += __shfl_xor_sync(0xffffffffu, sq[rr], off);
#pragma unroll
        for (int rr = 0; rr < ROWS; ++rr) {
            const int r = warp * ROWS + rr;
            float* row = sX + r * Cfg::X_PITCH;
            const float mean = sum[rr];
            const float rstd = rsqrtf(sq[rr] / (float)C + eps);
#pragma unroll
            for (int i = 0; i < PER / 4; ++i) {
                const int c4 = lane + 32 * i;
                float4 o;
                o.x = (v[rr][4 * i] - mean) * rstd * gm[i].x + bt[i].x;
                o.y = (v[rr][4 * i + 1] - mean) * rstd * gm[i].y + bt[i].y;
                o.z = (v[rr][4 * i + 2] - mean) * rstd * gm[i].z + bt[i].z;
                o.w = (v[rr][4 * i + 3] - mean) * rstd * gm[i].w + bt[i].w;
                if (to_global) {
                    if (row0 + r < p.T) *reinterpret_cast<float4*>(p.out + (size_t)(row0 + r) * p.ld_out + c4 * 4) = o;
                } else {
                    *reinterpret_cast<float4*>(row + c4 * 4) = o;  // x1 stays in fp32 for the second residual
                    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(sA + r * Cfg::A_PITCH + c4 * 8), "r"(dt_pack(o.x, o.y)),
                                 "r"(dt_pack(o.z, o.w)) : "memory");  // ... and as the expand GEMM's A operand
                }
            }
        }
    };

    // ---- GEMM 0: linear(hidden); + bias + residual -> sX; LayerNorm -> x1 (sX fp32, sA half) ----
    LnW ln1, ln2;
    {
        ln_load(p.g1, p.be1, ln1);
        float acc[2][C / 64][4] = {};
        float2 bias_r[C / 64];  // fetched before the GEMM: their L2 latency used to sit in front of every epilogue
#pragma unroll
        for (int nt = 0; nt < C / 64; ++nt) bias_r[nt] = __ldg(reinterpret_cast<const float2*>(p.b_lin + warp * (C / 8) + nt * 8 + 2 * q));
        dt_gemm<C / 64, 2>(acc, ring, 0, Cfg::CH0, sA, Cfg::A_PITCH, warp * (C / 8), lane, C * DT_ROW_BYTES);
        PA_MARK(2);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < C / 64; ++nt) {
                const int col = warp * (C / 8) + nt * 8 + 2 * q;
                const float2 bias = bias_r[nt];
#pragma unroll
                for (int h = 0; h < 2; ++h) {  // in place: each element of sX is read and written by one thread
                    float2* x = reinterpret_cast<float2*>(sX + (mt * 16 + g + 8 * h) * Cfg::X_PITCH + col);
                    const float2 res = *x;
                    *x = make_float2(acc[mt][nt][2 * h] + bias.x + res.x, acc[mt][nt][2 * h + 1] + bias.y + res.y);
                }
            }
    }
    dt_sync_compute();  // all of sX written; every warp is done reading sA (hidden)
    PA_MARK(3);
    layer_norm(ln1, p.eps1, false);
    dt_sync_compute();  // x1 complete in sA / sX
    PA_MARK(4);
    // ---- GEMM 1: ReLU(expand(x1) + bias) -> sE (half) ----
    {
        float acc[2][C / 32][4] = {};
        float2 bias_r[C / 32];
#pragma unroll
        for (int nt = 0; nt < C / 32; ++nt) bias_r[nt] = __ldg(reinterpret_cast<const float2*>(p.b_exp + warp * (C / 4) + nt * 8 + 2 * q));
        dt_gemm<C / 32>(acc, ring, Cfg::CH0, Cfg::CH1, sA, Cfg::A_PITCH, warp * (C / 4), lane);
        PA_MARK(5);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < C / 32; ++nt) {
                const int col = warp * (C / 4) + nt * 8 + 2 * q;
                const float2 bias = bias_r[nt];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int r = mt * 16 + g + 8 * h;
                    asm volatile("st.shared.b32 [%0], %1;" ::"r"(sE + r * Cfg::E_PITCH + col * 2),
                                 "r"(dt_pack(fmaxf(acc[mt][nt][2 * h] + bias.x, 0.f), fmaxf(acc[mt][nt][2 * h + 1] + bias.y, 0.f)))
                                 : "memory");
                }
            }
    }
    dt_sync_compute();  // sE complete
    PA_MARK(6);
    // ---- GEMM 2: squeeze(e) + bias + x1 -> sX (in place: each element is read and written by one thread); LayerNorm ----
    {
        ln_load(p.g2, p.be2, ln2);
        float acc[2][C / 64][4] = {};
        float2 bias_r[C / 64];
#pragma unroll
        for (int nt = 0; nt < C / 64; ++nt) bias_r[nt] = __ldg(reinterpret_cast<const float2*>(p.b_sq + warp * (C / 8) + nt * 8 + 2 * q));
        dt_gemm<C / 64, 2>(acc, ring, Cfg::CH0 + Cfg::CH1, Cfg::CH2, sE, Cfg::E_PITCH, warp * (C / 8), lane, C * DT_ROW_BYTES);
        PA_MARK(7);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < C / 64; ++nt) {
                const int col = warp * (C / 8) + nt * 8 + 2 * q;
                const float2 bias = bias_r[nt];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    float2* x = reinterpret_cast<float2*>(sX + (mt * 16 + g + 8 * h) * Cfg::X_PITCH + col);
                    const float2 x1 = *x;
                    *x = make_float2(acc[mt][nt][2 * h] + bias.x + x1.x, acc[mt][nt][2 * h + 1] + bias.y + x1.y);
                }
            }
    }
    dt_sync_compute();
    PA_MARK(8);
    layer_norm(ln2, p.eps2, true);
    PA_MARK(9);
}

__global__ void f32_to_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, long long count) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) dst[i] = __float2half_rn(src[i]);
}

// W [N][K] fp32 -> the packed half image of dense_tile.cuh; one thread per 16-byte unit
__global__ void pack_linear_weight_kernel(const float* __restrict__ w, int N, int K, uint4* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // = (c * N + r) * 4 + u
    if (i >= (long long)N * K / 8) return;
    const int u = (int)(i & 3);
    const long long cr = i >> 2;
    const int r = (int)(cr % N), c = (int)(cr / N);
    const float* src = w + (size_t)r * K + c * DT_KC + u * 8;
    const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
    uint4 v;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(v.x) : "f"(a.y), "f"(a.x));
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(v.y) : "f"(a.w), "f"(a.z));
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(v.z) : "f"(b.y), "f"(b.x));
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(v.w) : "f"(b.w), "f"(b.z));
    out[cr * 4 + (u ^ ((r >> 1) & 3))] = v;
}

}  // namespace geotx

using namespace geotx;

#ifdef PA_PHASE_CLOCKS
extern "C" int geotx_debug_k4f_phases(long long* host16) {
    return check_cuda(cudaMemcpyFromSymbol(host16, pa_phase_clock, sizeof(long long) * 16), "geotx_debug_k4f_phases");
}
#endif

extern "C" int geotx_f32_to_f16(const float* src, void* dst, int64_t count, void* stream) {
    GEOTX_REQUIRE(src && dst && count > 0, "geotx_f32_to_f16: bad arguments");
    const long long blocks = (count + 255) / 256;
    GEOTX_REQUIRE(blocks < 2147483647LL, "geotx_f32_to_f16: too many elements");
    f32_to_f16_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, (__half*)dst, count);
    return launch_status("geotx_f32_to_f16");
}

extern "C" int geotx_pack_linear_weight(const float* w, int N, int K, void* packed_f16, void* stream) {
    GEOTX_REQUIRE(w && packed_f16, "geotx_pack_linear_weight: null pointer");
    GEOTX_REQUIRE(N >= 8 && N % 8 == 0 && K >= DT_KC && K % DT_KC == 0,
                  "geotx_pack_linear_weight: N=%d must be a multiple of 8 and K=%d a multiple of %d", N, K, DT_KC);
    GEOTX_REQUIRE((((uintptr_t)w | (uintptr_t)packed_f16) & 15) == 0, "geotx_pack_linear_weight: pointers must be 16-byte aligned");
    const long long units = (long long)N * K / 8;
    pack_linear_weight_kernel<<<(unsigned)((units + 255) / 256), 256, 0, (cudaStream_t)stream>>>(w, N, K, (uint4*)packed_f16);
    return launch_status("geotx_pack_linear_weight");
}

extern "C" int geotx_post_attention(const float* hidden, int64_t ld_hidden, const float* x_in, int64_t ld_x, int64_t T,
                                    int C, const void* w_lin_f16, const float* b_lin, const float* gamma1,
                                    const float* beta1, float eps1, const void* w_expand_f16, const float* b_expand,
                                    const void* w_squeeze_f16, const float* b_squeeze, const float* gamma2,
                                    const float* beta2, float eps2, float* out, int64_t ld_out, void* stream) {
    GEOTX_REQUIRE(hidden && x_in && w_lin_f16 && b_lin && gamma1 && beta1 && w_expand_f16 && b_expand && w_squeeze_f16 &&
                      b_squeeze && gamma2 && beta2 && out, "geotx_post_attention: null pointer");
    GEOTX_REQUIRE(C == 128 || C == 256, "geotx_post_attention: C must be 128 or 256 (got %d)", C);
    GEOTX_REQUIRE(T > 0 && T < 2147483647LL / 2, "geotx_post_attention: bad token count");
    GEOTX_REQUIRE(ld_hidden >= C && ld_x >= C && ld_out >= C && ld_hidden % 4 == 0 && ld_x % 4 == 0 && ld_out % 4 == 0,
                  "geotx_post_attention: row strides must be >= C and multiples of 4 floats");
    GEOTX_REQUIRE((((uintptr_t)hidden | (uintptr_t)x_in | (uintptr_t)out | (uintptr_t)w_lin_f16 | (uintptr_t)w_expand_f16 |
                    (uintptr_t)w_squeeze_f16 | (uintptr_t)gamma1 | (uintptr_t)beta1 | (uintptr_t)gamma2 | (uintptr_t)beta2 |
                    (uintptr_t)b_lin | (uintptr_t)b_expand | (uintptr_t)b_squeeze) & 15) == 0,
                  "geotx_post_attention: pointers must be 16-byte aligned");
    PostAttnArgs p;
    p.hidden = hidden; p.ld_hidden = ld_hidden;
    p.x_in = x_in; p.ld_x = ld_x;
    p.out = out; p.ld_out = ld_out;
    p.w_lin = (const __half*)w_lin_f16; p.b_lin = b_lin;
    p.g1 = gamma1; p.be1 = beta1;
    p.w_exp = (const __half*)w_expand_f16; p.b_exp = b_expand;
    p.w_sq = (const __half*)w_squeeze_f16; p.b_sq = b_squeeze;
    p.g2 = gamma2; p.be2 = beta2;
    p.eps1 = eps1; p.eps2 = eps2;
    p.T = (int)T;
    const int grid = (int)((T + DT_BM - 1) / DT_BM);
    auto launch = [&](auto kernel, int smem) -> int {
        int rc = check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                            "geotx_post_attention: shared memory");
        if (rc) return rc;
        kernel<<<grid, DT_THREADS, smem, (cudaStream_t)stream>>>(p);
        return 0;
    };
    int rc = C == 256 ? launch(post_attention_kernel<256>, PACfg<256>::SMEM) : launch(post_attention_kernel<128>, PACfg<128>::SMEM);
    if (rc) return rc;
    return launch_status("geotx_post_attention");
}
