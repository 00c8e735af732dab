// K4: fused residual add + LayerNorm over the channel dimension.
//
// Replaces the `self.norm(hidden_states + input_states)` pattern of
// /root/reference/geotransformer/modules/transformer/rpe_transformer.py:103, vanilla_transformer.py:101 and
// output_layer.py:20 (torch: one elementwise add kernel + one LayerNorm kernel, two extra passes over the tokens).
// One warp per token: the row is read once with 16-byte loads, mean and biased variance (torch.nn.LayerNorm, eps inside
// the square root) are reduced with warp shuffles in two passes over registers, the result is written once.
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {

constexpr int LN_WARPS = 8;
constexpr int LN_MAX_VEC = 8;  // float4 per lane: C <= 1024

__global__ void __launch_bounds__(LN_WARPS * 32)
add_layernorm_kernel(const float* __restrict__ x, const float* __restrict__ res, const float* __restrict__ gamma,
                     const float* __restrict__ beta, long long rows, int C, float eps, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long row = (long long)blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int nvec = C >> 2;  // float4 per row
    const float4* xr = (const float4*)(x + row * C);
    const float4* rr = res ? (const float4*)(res + row * C) : nullptr;
    float4 v[LN_MAX_VEC];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAX_VEC; ++i) {
        const int c4 = lane + 32 * i;
        if (c4 < nvec) {
            float4 a = __ldg(xr + c4);
            if (rr) {
                const float4 b = __ldg(rr + c4);
                a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
            }
            v[i] = a;
            sum += (a.x + a.y) + (a.z + a.w);
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
    const float mean = sum / (float)C;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAX_VEC; ++i) {
        const int c4 = lane + 32 * i;
        if (c4 < nvec) {
            const float dx = v[i].x - mean, dy = v[i].y - mean, dz = v[i].z - mean, dw = v[i].w - mean;
            sq += (dx * dx + dy * dy) + (dz * dz + dw * dw);
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, off);
    const float rstd = rsqrtf(sq / (float)C + eps);
    float4* orow = (float4*)(out + row * C);
#pragma unroll
    for (int i = 0; i < LN_MAX_VEC; ++i) {
        const int c4 = lane + 32 * i;
        if (c4 < nvec) {
            const float4 g = __ldg((const float4*)gamma + c4), b = __ldg((const float4*)beta + c4);
            float4 o;
            o.x = (v[i].x - mean) * rstd * g.x + b.x;
            o.y = (v[i].y - mean) * rstd * g.y + b.y;
            o.z = (v[i].z - mean) * rstd * g.z + b.z;
            o.w = (v[i].w - mean) * rstd * g.w + b.w;
            orow[c4] = o;
        }
    }
}

}  // namespace geotx

using namespace geotx;

extern "C" int geotx_add_layernorm(const float* x, const float* residual, const float* gamma, const float* beta,
                                   int64_t rows, int C, float eps, float* out, void* stream) {
    GEOTX_REQUIRE(x && gamma && beta && out, "geotx_add_layernorm: null pointer");
    GEOTX_REQUIRE(rows > 0 && C >= 4 && C % 4 == 0 && C <= 128 * LN_MAX_VEC,
                  "geotx_add_layernorm: channels %d must be a multiple of 4 in [4, %d]", C, 128 * LN_MAX_VEC);
    GEOTX_REQUIRE((((uintptr_t)x | (uintptr_t)residual | (uintptr_t)gamma | (uintptr_t)beta | (uintptr_t)out) & 15) == 0,
                  "geotx_add_layernorm: pointers must be 16-byte aligned");
    const long long blocks = (rows + LN_WARPS - 1) / LN_WARPS;
    GEOTX_REQUIRE(blocks < 2147483647LL, "geotx_add_layernorm: too many rows");
    add_layernorm_kernel<<<(unsigned)blocks, LN_WARPS * 32, 0, (cudaStream_t)stream>>>(x, residual, gamma, beta, rows, C,
                                                                                        eps, out);
    return launch_status("geotx_add_layernorm");
}
