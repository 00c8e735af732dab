// K13: the token-wise Linear layers of the path (y = x W^T + b) that are not part of K4f's chain:
//   in_proj / out_proj                      geotransformer.py:140-141,152-153
//   q | k | v | folded-query projections    rpe_transformer.py:51-53 (+ proj_p folded, see geotx_rpe_scores),
//                                           vanilla_transformer.py:50-52; several Linears of one block concatenated
//                                           along the output dimension into one call
// Round 1 issued these through cuBLASLt (TF32).  At ~2k tokens each call is a few microseconds of work, so what
// matters is that it is one short launch with the bias in the epilogue and no layout round trip: x is read as fp32
// and rounded to half on its way into shared memory (same 11-bit significand as TF32), W comes from a half copy that
// is made once per weight set ([N][K], nn.Linear's own layout = the mma "col" operand), cp.async double buffering,
// mma.sync.m16n8k16 with fp32 accumulation, fp32 output with arbitrary row strides on both sides (so that inputs and
// outputs can be column blocks / row slices of larger token buffers).
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {

constexpr int LIN_BN = 64, LIN_KC = 32, LIN_THREADS = 128;
constexpr int LIN_PITCH = LIN_KC * 2 + 16;  // 80 B per row of a stage: odd multiple of 16 (conflict-free ldmatrix)

__device__ __forceinline__ void lin_ldsm4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void lin_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t lin_pack(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// CTA tile BM x 64, 4 warps as 2 x 2 (warp tile BM/2 x 32).  BM = 64 or 32 (small grids).
template <int BM>
__global__ void __launch_bounds__(LIN_THREADS)
linear_f16_kernel(const float* __restrict__ x, long long ld_x, int T, int K, const __half* __restrict__ w,
                  const float* __restrict__ bias, int N, float* __restrict__ y, long long ld_y, int relu) {
    constexpr int MT = BM / 32;                      // m-tiles per warp
    constexpr int A_F4 = BM * (LIN_KC / 4) / LIN_THREADS;  // float4 loads of x per thread per chunk (4 or 2)
    __shared__ __align__(16) unsigned char smem[2 * (BM + LIN_BN) * LIN_PITCH];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 1, wn = warp & 1;
    const int row0 = blockIdx.y * BM, col0 = blockIdx.x * LIN_BN;
    const uint32_t sA = smem_u32(smem), sB = sA + 2 * BM * LIN_PITCH;

    float4 areg[A_F4];
    auto load_a = [&](int chunk) {
#pragma unroll
        for (int i = 0; i < A_F4; ++i) {
            const int e = tid + i * LIN_THREADS, r = e >> 3, c4 = e & 7;
            areg[i] = row0 + r < T ? __ldg(reinterpret_cast<const float4*>(x + (size_t)(row0 + r) * ld_x + chunk * LIN_KC) + c4)
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };
    auto store_a = [&](int stage) {
#pragma unroll
        for (int i = 0; i < A_F4; ++i) {
            const int e = tid + i * LIN_THREADS, r = e >> 3, c4 = e & 7;
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(sA + (stage * BM + r) * LIN_PITCH + c4 * 8),
                         "r"(lin_pack(areg[i].x, areg[i].y)), "r"(lin_pack(areg[i].z, areg[i].w)) : "memory");
        }
    };
    auto load_b = [&](int chunk, int stage) {
#pragma unroll
        for (int i = 0; i < LIN_BN * 4 / LIN_THREADS; ++i) {
            const int e = tid + i * LIN_THREADS, n = e >> 2, part = e & 3;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sB + (stage * LIN_BN + n) * LIN_PITCH + part * 16),
                         "l"(w + (size_t)(col0 + n) * K + chunk * LIN_KC + part * 8) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    float acc[MT][4][4] = {};
    const int chunks = K / LIN_KC;
    load_a(0);
    load_b(0, 0);
    store_a(0);
    for (int c = 0; c < chunks; ++c) {
        const int st = c & 1;
        if (c + 1 < chunks) {
            load_a(c + 1);          // global -> registers, consumed after this chunk's math
            load_b(c + 1, st ^ 1);  // stage st^1 was last read in iteration c-1 (barrier at its end)
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();  // chunk c's A (st.shared) and B (cp.async) are visible to every warp
#pragma unroll
        for (int ks = 0; ks < LIN_KC / 16; ++ks) {
            uint32_t a[MT][4];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
                lin_ldsm4(a[mt], sA + (st * BM + wm * (BM / 2) + mt * 16 + (lane & 15)) * LIN_PITCH + (ks * 16 + (lane >> 4) * 8) * 2);
#pragma unroll
            for (int np = 0; np < 2; ++np) {
                uint32_t b[4];
                lin_ldsm4(b, sB + (st * LIN_BN + wn * 32 + np * 16 + (lane & 7) + ((lane >> 4) & 1) * 8) * LIN_PITCH +
                                 (ks * 16 + ((lane >> 3) & 1) * 8) * 2);
#pragma unroll
                for (int mt = 0; mt < MT; ++mt) {
                    lin_mma(acc[mt][2 * np], a[mt], b[0], b[1]);
                    lin_mma(acc[mt][2 * np + 1], a[mt], b[2], b[3]);
                }
            }
        }
        if (c + 1 < chunks) store_a(st ^ 1);
        __syncthreads();  // everyone is done with stage st before iteration c+1 refills it (and sees A of c+1)
    }
    const int g = lane >> 2, q = lane & 3;
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
            const int col = col0 + wn * 32 + nt * 8 + 2 * q;
            const float2 b = bias ? __ldg(reinterpret_cast<const float2*>(bias + col)) : make_float2(0.f, 0.f);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int r = row0 + wm * (BM / 2) + mt * 16 + g + 8 * h;
                if (r < T) {
                    float2 o = make_float2(acc[mt][nt][2 * h] + b.x, acc[mt][nt][2 * h + 1] + b.y);
                    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); }
                    *reinterpret_cast<float2*>(y + (size_t)r * ld_y + col) = o;
                }
            }
        }
}

}  // namespace geotx

using namespace geotx;

extern "C" int geotx_linear_f16(const float* x, int64_t ld_x, int64_t T, int K, const void* w_f16, const float* bias,
                                int N, int relu, float* y, int64_t ld_y, void* stream) {
    GEOTX_REQUIRE(x && w_f16 && y, "geotx_linear_f16: null pointer");
    GEOTX_REQUIRE(T > 0 && T < 2147483647LL / 2, "geotx_linear_f16: bad token count");
    GEOTX_REQUIRE(K >= LIN_KC && K % LIN_KC == 0 && N >= LIN_BN && N % LIN_BN == 0,
                  "geotx_linear_f16: K=%d must be a multiple of %d and N=%d a multiple of %d", K, LIN_KC, N, LIN_BN);
    GEOTX_REQUIRE(ld_x >= K && ld_x % 4 == 0 && ld_y >= N && ld_y % 2 == 0,
                  "geotx_linear_f16: row strides (x: multiple of 4 floats >= K, y: even >= N)");
    GEOTX_REQUIRE((((uintptr_t)x | (uintptr_t)w_f16) & 15) == 0 && (((uintptr_t)y | (uintptr_t)bias) & 7) == 0,
                  "geotx_linear_f16: x, w must be 16-byte aligned, y and bias 8-byte aligned");
    int dev = 0, sms = 0;
    int rc = check_cuda(cudaGetDevice(&dev), "geotx_linear_f16: cudaGetDevice");
    if (rc) return rc;
    rc = check_cuda(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev), "geotx_linear_f16: SM count");
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const long long tiles64 = ((T + 63) / 64) * (N / LIN_BN);
    if (tiles64 >= 2LL * sms) {
        dim3 grid(N / LIN_BN, (unsigned)((T + 63) / 64));
        linear_f16_kernel<64><<<grid, LIN_THREADS, 0, st>>>(x, ld_x, (int)T, K, (const __half*)w_f16, bias, N, y, ld_y, relu);
    } else {  // few tiles: halve them so that more SMs take part
        dim3 grid(N / LIN_BN, (unsigned)((T + 31) / 32));
        linear_f16_kernel<32><<<grid, LIN_THREADS, 0, st>>>(x, ld_x, (int)T, K, (const __half*)w_f16, bias, N, y, ld_y, relu);
    }
    return launch_status("geotx_linear_f16");
}
