// K13: the token-wise Linear layers of the path (y = x W^T + b) that are not part of K4f's chain:
//   in_proj / out_proj                      geotransformer.py:140-141,152-153
//   q | k | v | folded-query projections    rpe_transformer.py:51-53 (+ proj_p folded, see geotx_rpe_scores),
//                                           vanilla_transformer.py:50-52; several Linears of one block concatenated
//                                           along the output dimension into one call
// Round 1 issued these through cuBLASLt (TF32).  Same building block as K4f (dense_tile.cuh): a CTA owns 32 tokens and
// up to 256 output columns; x is read ONCE as fp32 and rounded to half into shared memory (same 11-bit significand as
// TF32), the packed weights (geotx_pack_linear_weight) stream through a 4-stage ring of TMA bulk copies issued by a
// loader warp, eight compute warps contract on mma.sync.m16n8k16 with fp32 accumulation and write fp32 with the bias
// (and ReLU) in the epilogue; arbitrary row strides on both sides, so inputs and outputs can be column blocks / row
// slices of larger token buffers.  The first version (64 x 64 tiles, 16-byte cp.async from the compute warps, a
// __syncthreads per K chunk, x re-converted by every column tile) took 0.9 us per K chunk: 28 us for in_proj (K = 1024).
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "host_util.h"
#include "sm100_ptx.cuh"
#include "dense_tile.cuh"

namespace geotx {

struct LinearArgs {
    const float* x; long long ld_x;
    const __half* w;     // packed [K/32][N][32]
    const float* bias;   // [N] or null
    float* y; long long ld_y;
    int T, K, N, relu;
    __half* yh; long long ld_yh;  // geotx_linear_f16_mixed: output columns < half_cols go here as half ...
    int half_cols;                // ... and column c >= half_cols to y[.., c - half_cols] (0: everything fp32 into y)
};

constexpr int LIN_OFF_A = 128;
__host__ __device__ constexpr int lin_a_pitch(int K) { return K * 2 + 16; }
__host__ __device__ constexpr int lin_off_w(int K) { return (LIN_OFF_A + DT_BM * lin_a_pitch(K) + 127) / 128 * 128; }

// NT = n-tiles per warp: the CTA covers NB = 64 NT output columns from blockIdx.x * NB (fewer in the last CTA of a row)
template <int NT>
__global__ void __launch_bounds__(DT_THREADS) linear_tile_kernel(const LinearArgs p) {
    constexpr int NB = 64 * NT;
    extern __shared__ __align__(128) unsigned char lin_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int row0 = blockIdx.y * DT_BM, n0 = blockIdx.x * NB;
    const int cols = min(NB, p.N - n0);  // a multiple of 64
    const uint32_t s_base = smem_u32(lin_smem);
    const uint32_t sA = s_base + LIN_OFF_A;
    const int a_pitch = lin_a_pitch(p.K);
    DtRing ring;
    ring.bar_full = s_base;
    ring.bar_empty = s_base + 8 * DT_STAGES;
    ring.stages = s_base + lin_off_w(p.K);
    ring.stage_bytes = NB * DT_ROW_BYTES;
    ring.init(tid);
    __syncthreads();
    const int chunks = p.K / DT_KC;

    if (warp == DT_WARPS) {  // ---- loader ----
        if (lane == 0) {
            for (int c = 0; c < chunks; ++c) ring.load(c, p.w + ((size_t)c * p.N + n0) * DT_KC, cols * DT_ROW_BYTES);
#ifdef DT_LOADER_STAYS
            for (int c = max(chunks - DT_STAGES, 0); c < chunks; ++c) ring.wait(c);
#endif
        }
        return;
    }

    // the tile's tokens -> half A operand, once.  Sixteen independent 16-byte loads in flight per thread: at K = 1024 the
    // tile is 128 KB of fp32 from L2 / HBM, and a loop of dependent load -> convert -> store steps is latency-bound
    const int k4 = p.K / 4, total4 = DT_BM * k4;
    constexpr int U = 16;  // loads in flight per thread (8: in_proj's tile, 32 float4 per thread, paid four round trips)
    for (int e0 = tid; e0 < total4; e0 += U * DT_WARPS * 32) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int e = e0 + u * DT_WARPS * 32, r = e / k4, c4 = e - r * k4;
            v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (e < total4 && row0 + r < p.T) v[u] = __ldg(reinterpret_cast<const float4*>(p.x + (size_t)(row0 + r) * p.ld_x) + c4);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int e = e0 + u * DT_WARPS * 32, r = e / k4, c4 = e - r * k4;
            if (e < total4)
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(sA + r * a_pitch + c4 * 8), "r"(dt_pack(v[u].x, v[u].y)),
                             "r"(dt_pack(v[u].z, v[u].w)) : "memory");
        }
    }
    dt_sync_compute();

    const int n_warp = warp * (8 * NT);  // this warp's first column inside the CTA's block
    if (n_warp >= cols) {                // ragged last column block: nothing to compute, but the ring counts 8 warps
        for (int c = 0; c < chunks; ++c) {
            ring.wait(c);
            ring.release(c, lane);
        }
        return;
    }
    float acc[2][NT][4] = {};
    const int g = lane >> 2, q = lane & 3;
    float2 bias_r[NT];  // fetched before the GEMM (their L2 latency used to sit in front of the epilogue)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
        bias_r[nt] = p.bias ? __ldg(reinterpret_cast<const float2*>(p.bias + n0 + n_warp + nt * 8 + 2 * q)) : make_float2(0.f, 0.f);
    dt_gemm<NT>(acc, ring, 0, chunks, sA, a_pitch, n_warp, lane);
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        const int col = n0 + n_warp + nt * 8 + 2 * q;
        const float2 b = bias_r[nt];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int r = row0 + mt * 16 + g + 8 * h;
                if (r < p.T) {
                    float2 o = make_float2(acc[mt][nt][2 * h] + b.x, acc[mt][nt][2 * h + 1] + b.y);
                    if (p.relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); }
                    if (col < p.half_cols)  // uniform per 64-column block: half_cols is a multiple of 64
                        *reinterpret_cast<uint32_t*>(p.yh + (size_t)r * p.ld_yh + col) = dt_pack(o.x, o.y);
                    else
                        *reinterpret_cast<float2*>(p.y + (size_t)r * p.ld_y + (col - p.half_cols)) = o;
                }
            }
    }
}

}  // namespace geotx

using namespace geotx;

static int linear_launch(const float* x, int64_t ld_x, int64_t T, int K, const void* w_f16, const float* bias, int N, int relu,
                         float* y, int64_t ld_y, void* y_half, int64_t ld_yh, int half_cols, void* stream);

extern "C" int geotx_linear_f16(const float* x, int64_t ld_x, int64_t T, int K, const void* w_f16, const float* bias,
                                int N, int relu, float* y, int64_t ld_y, void* stream) {
    GEOTX_REQUIRE(y, "geotx_linear_f16: null pointer");
    return linear_launch(x, ld_x, T, K, w_f16, bias, N, relu, y, ld_y, nullptr, 0, 0, stream);
}

extern "C" int geotx_linear_f16_mixed(const float* x, int64_t ld_x, int64_t T, int K, const void* w_f16, const float* bias,
                                      int N, int relu, void* y_half, int64_t ld_yh, int half_cols, float* y, int64_t ld_y,
                                      void* stream) {
    GEOTX_REQUIRE(half_cols > 0 && half_cols <= N && half_cols % 64 == 0,
                  "geotx_linear_f16_mixed: half_cols=%d must be a multiple of 64 in (0, N=%d]", half_cols, N);
    GEOTX_REQUIRE(y_half && ((uintptr_t)y_half & 15) == 0 && ld_yh >= half_cols && ld_yh % 8 == 0,
                  "geotx_linear_f16_mixed: y_half must be 16-byte aligned with a row stride >= half_cols, multiple of 8 halves");
    GEOTX_REQUIRE(half_cols == N || y, "geotx_linear_f16_mixed: y is null but %d fp32 columns remain", N - half_cols);
    return linear_launch(x, ld_x, T, K, w_f16, bias, N, relu, y, ld_y, y_half, ld_yh, half_cols, stream);
}

static int linear_launch(const float* x, int64_t ld_x, int64_t T, int K, const void* w_f16, const float* bias, int N, int relu,
                         float* y, int64_t ld_y, void* y_half, int64_t ld_yh, int half_cols, void* stream) {
    GEOTX_REQUIRE(x && w_f16, "geotx_linear_f16: null pointer");
    GEOTX_REQUIRE(T > 0 && T < 2147483647LL / 2, "geotx_linear_f16: bad token count");
    GEOTX_REQUIRE(K >= DT_KC && K % DT_KC == 0 && K <= 2048 && N >= 64 && N % 64 == 0,
                  "geotx_linear_f16: K=%d must be a multiple of %d (<= 2048) and N=%d a multiple of 64", K, DT_KC, N);
    GEOTX_REQUIRE(ld_x >= K && ld_x % 4 == 0 && (half_cols == N || (ld_y >= N - half_cols && ld_y % 2 == 0)),
                  "geotx_linear_f16: row strides (x: multiple of 4 floats >= K, y: even >= its column count)");
    GEOTX_REQUIRE((((uintptr_t)x | (uintptr_t)w_f16) & 15) == 0 && (((uintptr_t)y | (uintptr_t)bias) & 7) == 0,
                  "geotx_linear_f16: x, w must be 16-byte aligned, y and bias 8-byte aligned");
    LinearArgs p;
    p.x = x; p.ld_x = ld_x;
    p.w = (const __half*)w_f16;
    p.bias = bias;
    p.y = y; p.ld_y = ld_y;
    p.T = (int)T; p.K = K; p.N = N; p.relu = relu;
    p.yh = (__half*)y_half; p.ld_yh = ld_yh; p.half_cols = half_cols;
    const int sms = sm_count();
    const long long row_tiles = (T + DT_BM - 1) / DT_BM;
    // the widest column block that still gives every SM a CTA (x is converted once per CTA: wider is cheaper)
    int nt = 4;
    while (nt > 1 && row_tiles * ((N + 64 * nt - 1) / (64 * nt)) < sms) nt >>= 1;
#ifdef LIN_SMALL_NT  // timing experiment: column-block width when no width fills the GPU
    if (row_tiles * ((N + 63) / 64) < sms) nt = N % (64 * LIN_SMALL_NT) == 0 ? LIN_SMALL_NT : 1;
#endif
#ifdef DT_ONE_CTA
    const size_t smem = 120 * 1024;
#else
    const size_t smem = (size_t)lin_off_w(K) + (size_t)DT_STAGES * 64 * nt * DT_ROW_BYTES;
#endif
    const dim3 grid((unsigned)((N + 64 * nt - 1) / (64 * nt)), (unsigned)row_tiles);
    auto launch = [&](auto kernel) -> int {
        int rc = check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "geotx_linear_f16: shared memory");
        if (rc) return rc;
        kernel<<<grid, DT_THREADS, smem, (cudaStream_t)stream>>>(p);
        return 0;
    };
    int rc = nt == 4 ? launch(linear_tile_kernel<4>) : nt == 2 ? launch(linear_tile_kernel<2>) : launch(linear_tile_kernel<1>);
    if (rc) return rc;
    return launch_status("geotx_linear_f16");
}
