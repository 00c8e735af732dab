// K2 (fp32 variant): GeometricStructureEmbedding.forward on CUDA cores, fp32 throughout, accurate sinf/cosf.
// Replaces /root/reference/geotransformer/modules/geotransformer/geotransformer.py:57-72 and
// /root/reference/geotransformer/modules/transformer/positional_embedding.py:18-34 for ANY even hidden size.
// This is the high-precision path (small hidden sizes, on-device cross-check of the tcgen05 kernel); the
// throughput path is k2_gse_embed.cu.  Neither ever materialises the reference's (B,n,n,k,C) tensor.
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {

constexpr int F32_PAIRS = 8;  // pairs per CTA

// grid.x = pair tiles of one slab, block = C threads (thread c owns output channel c).
// smem: xs[F32_PAIRS][8] | sv[F32_PAIRS][k+1][C]
__global__ void gse_embed_fp32_kernel(const float4* __restrict__ pts4, const float4* __restrict__ refv, int n, int k,
                                      int row_begin, int row_count, long long pairs_total,
                                      const float* __restrict__ div_term, const float* __restrict__ w_d,
                                      const float* __restrict__ b_d, const float* __restrict__ w_a,
                                      const float* __restrict__ b_a, int C, float sigma_d, float factor_a,
                                      int reduce_mean, int out_dtype, void* __restrict__ out) {
    extern __shared__ float smem[];
    float* xs = smem;                    // [F32_PAIRS][8]
    float* sv = smem + F32_PAIRS * 8;    // [F32_PAIRS][k+1][C]
    const int c = threadIdx.x;
    const long long p0 = (long long)blockIdx.x * F32_PAIRS;
    const int slots = k + 1;
    const long long per_b = (long long)row_count * n;

    if (c < F32_PAIRS) {
        long long p = p0 + c;
        float x[GEOTX_MAX_K + 1];
#pragma unroll
        for (int q = 0; q <= GEOTX_MAX_K; ++q) x[q] = 0.f;
        if (p < pairs_total) {
            long long b = p / per_b;
            long long rem = p - b * per_b;
            int i = row_begin + (int)(rem / n);
            int j = (int)(rem % n);
            const float4 pi = __ldg(pts4 + b * n + i);
            const float4 pj = __ldg(pts4 + b * n + j);
            pair_arguments<GEOTX_MAX_K>(pi, pj, refv + (b * n + i) * k, k, sigma_d, factor_a, x);
        }
#pragma unroll
        for (int q = 0; q < GEOTX_MAX_K; ++q) xs[c * 8 + q] = x[q];
        xs[c * 8 + 7] = x[GEOTX_MAX_K];  // distance argument lives in column 7
    }
    __syncthreads();

    // positional_embedding.py:28-31: omega = x * div_term[t]; emb[2t] = sin(omega), emb[2t+1] = cos(omega)
    const float dv = div_term[c >> 1];
    for (int p = 0; p < F32_PAIRS; ++p) {
        for (int s = 0; s < slots; ++s) {
            float x = (s < k) ? xs[p * 8 + s] : xs[p * 8 + 7];
            float om = __fmul_rn(x, dv);
            sv[(p * slots + s) * C + c] = (c & 1) ? cosf(om) : sinf(om);
        }
    }
    __syncthreads();

    float acc_d[F32_PAIRS];
    float acc_a[F32_PAIRS][GEOTX_MAX_K];
#pragma unroll
    for (int p = 0; p < F32_PAIRS; ++p) {
        acc_d[p] = 0.f;
#pragma unroll
        for (int q = 0; q < GEOTX_MAX_K; ++q) acc_a[p][q] = 0.f;
    }
    const float* wd_row = w_d + (size_t)c * C;
    const float* wa_row = w_a + (size_t)c * C;
    for (int kc = 0; kc < C; ++kc) {
        const float wd = __ldg(wd_row + kc);
        const float wa = __ldg(wa_row + kc);
#pragma unroll
        for (int p = 0; p < F32_PAIRS; ++p) {
            const float* s = sv + (size_t)p * slots * C + kc;
#pragma unroll
            for (int q = 0; q < GEOTX_MAX_K; ++q)
                if (q < k) acc_a[p][q] = fmaf(wa, s[q * C], acc_a[p][q]);
            acc_d[p] = fmaf(wd, s[k * C], acc_d[p]);
        }
    }
    const float bd = b_d[c], ba = b_a[c];
#pragma unroll
    for (int p = 0; p < F32_PAIRS; ++p) {
        long long pp = p0 + p;
        if (pp >= pairs_total) break;
        float red;
        if (reduce_mean) {
            red = 0.f;
#pragma unroll
            for (int q = 0; q < GEOTX_MAX_K; ++q)
                if (q < k) red += acc_a[p][q] + ba;
            red = red / (float)k;
        } else {
            red = -INFINITY;
#pragma unroll
            for (int q = 0; q < GEOTX_MAX_K; ++q)
                if (q < k) red = fmaxf(red, acc_a[p][q] + ba);
        }
        float e = (acc_d[p] + bd) + red;  // geotransformer.py:70
        if (out_dtype == GEOTX_BF16) ((__nv_bfloat16*)out)[pp * C + c] = __float2bfloat16_rn(e);
        else if (out_dtype == GEOTX_F16) ((__half*)out)[pp * C + c] = __float2half_rn(e);
        else ((float*)out)[pp * C + c] = e;
    }
}

}  // namespace geotx

using namespace geotx;

extern "C" int geotx_gse_embed_fp32(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                                    const float* div_term, const float* w_d, const float* b_d, const float* w_a,
                                    const float* b_a, int C, float sigma_d, float factor_a, int reduce_mean,
                                    int out_dtype, void* out, void* stream) {
    GEOTX_REQUIRE(geometry && div_term && w_d && b_d && w_a && b_a && out, "geotx_gse_embed_fp32: null pointer");
    GEOTX_REQUIRE(B > 0 && n > 0 && k >= 1 && k <= GEOTX_MAX_ANGLE_K, "geotx_gse_embed_fp32: bad sizes");
    GEOTX_REQUIRE(C >= 2 && C % 2 == 0 && C <= 1024, "geotx_gse_embed_fp32: hidden_dim %d must be even and <= 1024", C);
    GEOTX_REQUIRE(row_begin >= 0 && row_count > 0 && row_begin + row_count <= n, "geotx_gse_embed_fp32: bad row slab");
    GEOTX_REQUIRE(out_dtype == GEOTX_F32 || out_dtype == GEOTX_BF16 || out_dtype == GEOTX_F16,
                  "geotx_gse_embed_fp32: bad out_dtype");
    const float4* pts4 = (const float4*)geometry;
    const float4* refv = pts4 + (size_t)B * n;
    long long pairs = (long long)B * row_count * n;
    long long blocks = (pairs + F32_PAIRS - 1) / F32_PAIRS;
    GEOTX_REQUIRE(blocks < 2147483647LL, "geotx_gse_embed_fp32: too many pairs");
    size_t smem = sizeof(float) * (F32_PAIRS * 8 + (size_t)F32_PAIRS * (k + 1) * C);
    GEOTX_REQUIRE(smem <= 200 * 1024, "geotx_gse_embed_fp32: hidden_dim*k too large for shared memory");
    int rc = check_cuda(cudaFuncSetAttribute(gse_embed_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem), "cudaFuncSetAttribute");
    if (rc) return rc;
    gse_embed_fp32_kernel<<<(unsigned)blocks, C, smem, (cudaStream_t)stream>>>(
        pts4, refv, n, k, row_begin, row_count, pairs, div_term, w_d, b_d, w_a, b_a, C, sigma_d, factor_a,
        reduce_mean, out_dtype, out);
    return launch_status("geotx_gse_embed_fp32");
}
