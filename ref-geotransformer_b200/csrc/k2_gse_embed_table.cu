// K2t (round 1's default embedding; since round 2 a variant, GEOTX_EMBED=table -- the default is the low-rank kernel
// k2_gse_embed_cheb.cu): GeometricStructureEmbedding by TABULATION instead of by contraction.
//
// geotransformer.py:57-72 computes E[i,j,:] = proj_d(sinus(d_ij)) + max_k proj_a(sinus(a_ijk)).  Per output channel,
// proj_x(sinus(x)) + bias is a function of ONE scalar x (positional_embedding.py:28-31 followed by a Linear):
//     f_c(x) = b[c] + sum_t W[c,2t] sin(x div_t) + W[c,2t+1] cos(x div_t),      div_t <= 1,
// smooth at the scale of one index unit (the highest frequency is 1 rad per unit).  K2 (k2_gse_embed.cu) spends
// 2 C^2 (1+k) flop per pair on it on the tensor cores; here f_c is tabulated once per weight set (tables fa, fd:
// rows x C halves, bias folded in) and interpolated out of shared memory: the angle term linearly on a fine grid
// (the step is the caller's: modules.py ships h_a = 1/8; two rows per slot, error <= h^2/8 max|f"|, i.e. it grows with
// the weights -- unlike the Chebyshev form, whose error is relative to them), the distance term, whose range is larger, by
// cubic Lagrange interpolation on a coarse grid (shipped step h_d = 1/2: four rows).  10 row reads (5 KB) and ~30
// packed half FMAs per pair for C = 256, k = 3: bound by the shared-memory pipe instead of the tensor pipe, and below
// the fp16 rounding of the stored E (and of K2's bf16 operands).
//
// One warp per pair: lane l owns C/32 consecutive channels (one LDS.128 per table row and one 16-byte store of E at
// C = 256, LDS.64 / 8 bytes at C = 128).  The lanes of
// a warp first compute the arguments of 32 consecutive pairs (one each), then walk them together; a lookup travels
// between lanes as one word.
#include <cuda_fp16.h>
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {

// E is written once and never read by this SM: no L1 line is allocated for it (the stores share the L1 / shared-memory
// path with the table reads).
__device__ __forceinline__ void tbl_store(void* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void tbl_store(void* p, const uint2& v) {
    asm volatile("st.global.L1::no_allocate.v2.b32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

template <int C> struct TblVec;  // what one lane reads of a table row: C / 32 halves
template <> struct TblVec<256> { typedef uint4 type; };
template <> struct TblVec<128> { typedef uint2 type; };
constexpr int TBL_THREADS = 512;
constexpr int TBL_CTAS_PER_SM = 2;  // tables up to ~110 KB

// Row r of a table holds f(x) at x = (r - first) * h (first = 1 for the cubic table, so that the four rows around any
// x >= 0 exist).
__global__ void gse_table_build_kernel(const float* __restrict__ div_term, const float* __restrict__ w,
                                       const float* __restrict__ bias, int C, float h, int first,
                                       __half* __restrict__ table) {
    const int r = blockIdx.x, c = threadIdx.x;
    const double x = (double)(r - first) * (double)h;
    double acc = (double)bias[c];
    for (int t = 0; t < C / 2; ++t) {
        const double om = x * (double)div_term[t];
        acc += (double)w[(size_t)c * C + 2 * t] * sin(om) + (double)w[(size_t)c * C + 2 * t + 1] * cos(om);
    }
    table[(size_t)r * C + c] = __float2half_rn((float)acc);
}

// A lookup: table row (high 16 bits) and the position inside the interval as a 16-bit fraction.  `span` rows are read
// (2 linear, 4 cubic).
__device__ __forceinline__ unsigned int tbl_encode(float x, float inv_h, int rows, int span) {
    const float u = fmaxf(x * inv_h, 0.f);
    const float fl = fminf(floorf(u), (float)(rows - span));  // arguments beyond the table are flagged by the caller
    const float t = fminf(u - fl, 0.99998f);
    return ((unsigned int)fl << 16) | (unsigned int)(t * 65536.f);
}

// C / 32 channels (C / 64 half2) of lane `lane`: linear interpolation between rows r and r + 1
template <int C>
__device__ __forceinline__ void tbl_linear(const __half* __restrict__ table, unsigned int code, int lane,
                                           __half2 (&acc)[C / 64]) {
    typedef typename TblVec<C>::type V;
    const __half2 t = __float2half2_rn((float)(code & 0xffffu) * (1.f / 65536.f));
    const V* base = reinterpret_cast<const V*>(table + (size_t)(code >> 16) * C) + lane;
    const V r0 = base[0], r1 = base[32];
    const __half2* p0 = reinterpret_cast<const __half2*>(&r0);
    const __half2* p1 = reinterpret_cast<const __half2*>(&r1);
#pragma unroll
    for (int q = 0; q < C / 64; ++q) acc[q] = __hfma2(t, __hsub2(p1[q], p0[q]), p0[q]);
}

// cubic Lagrange interpolation over rows r .. r + 3 (nodes at -1, 0, 1, 2 around the interval)
template <int C>
__device__ __forceinline__ void tbl_cubic(const __half* __restrict__ table, unsigned int code, int lane,
                                          __half2 (&acc)[C / 64]) {
    typedef typename TblVec<C>::type V;
    const float t = (float)(code & 0xffffu) * (1.f / 65536.f), tm1 = t - 1.f, tm2 = t - 2.f, tp1 = t + 1.f;
    const __half2 w0 = __float2half2_rn(-t * tm1 * tm2 * (1.f / 6.f)), w1 = __float2half2_rn(tp1 * tm1 * tm2 * 0.5f),
                  w2 = __float2half2_rn(-tp1 * t * tm2 * 0.5f), w3 = __float2half2_rn(tp1 * t * tm1 * (1.f / 6.f));
    const V* base = reinterpret_cast<const V*>(table + (size_t)(code >> 16) * C) + lane;
    const V r0 = base[0], r1 = base[32], r2 = base[64], r3 = base[96];
    const __half2* p0 = reinterpret_cast<const __half2*>(&r0);
    const __half2* p1 = reinterpret_cast<const __half2*>(&r1);
    const __half2* p2 = reinterpret_cast<const __half2*>(&r2);
    const __half2* p3 = reinterpret_cast<const __half2*>(&r3);
#pragma unroll
    for (int q = 0; q < C / 64; ++q)
        acc[q] = __hfma2(w3, p3[q], __hfma2(w2, p2[q], __hfma2(w1, p1[q], __hmul2(w0, p0[q]))));
}

// A distance beyond the table (a cloud larger than the range the tables were built for): the function itself,
// positional_embedding.py:28-31 + the Linear, for this lane's 8 channels.  Rare and slow (128 sincos + 2048 FMA).
template <int C>
__device__ __noinline__ void tbl_exact(const float* __restrict__ div_term, const float* __restrict__ w,
                                       const float* __restrict__ bias, float x, int lane, __half2 (&acc)[C / 64]) {
    constexpr int PER = C / 32;
    float a[PER];
#pragma unroll
    for (int q = 0; q < PER; ++q) a[q] = bias[PER * lane + q];
    for (int t = 0; t < C / 2; ++t) {
        float sn, cs;
        sincosf(__fmul_rn(x, div_term[t]), &sn, &cs);
#pragma unroll
        for (int q = 0; q < PER; ++q) {
            const float2 wt = *reinterpret_cast<const float2*>(w + (size_t)(PER * lane + q) * C + 2 * t);
            a[q] = fmaf(wt.y, cs, fmaf(wt.x, sn, a[q]));
        }
    }
#pragma unroll
    for (int q = 0; q < C / 64; ++q) acc[q] = __floats2half2_rn(a[2 * q], a[2 * q + 1]);
}

template <int K, int C>
__global__ void __launch_bounds__(TBL_THREADS, TBL_CTAS_PER_SM)
gse_embed_table_kernel(const float4* __restrict__ pts4, const float4* __restrict__ refv, int n, int row_begin,
                       int row_count, long long pairs_total, const __half* __restrict__ tables, int rows_a, int rows_d,
                       float inv_ha, float inv_hd, float sigma_d, float factor_a, const float* __restrict__ div_term,
                       const float* __restrict__ w_d, const float* __restrict__ b_d, __half* __restrict__ out,
                       int* __restrict__ overflow) {
    extern __shared__ __align__(16) unsigned char tbl_smem[];
    __half* fa = reinterpret_cast<__half*>(tbl_smem);
    typedef typename TblVec<C>::type V;
    __half* fd = fa + (size_t)rows_a * C;
    {
        const uint4* src = reinterpret_cast<const uint4*>(tables);
        uint4* dst = reinterpret_cast<uint4*>(tbl_smem);
        const int total = (rows_a + rows_d) * C / 8;
        for (int e = threadIdx.x; e < total; e += TBL_THREADS) dst[e] = src[e];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long warp = (long long)blockIdx.x * (TBL_THREADS / 32) + (threadIdx.x >> 5);
    const long long warps = (long long)gridDim.x * (TBL_THREADS / 32);
    const long long per_b = (long long)row_count * n;
    const float x_max_a = (float)(rows_a - 1) / inv_ha, x_max_d = (float)(rows_d - 3) / inv_hd;
    bool clamped = false;
    for (long long g0 = warp * 32; g0 < pairs_total; g0 += warps * 32) {
        unsigned int code[K + 1];  // this lane's pair of the group: arguments -> lookups
        float x_d;
        {
            const long long p = min(g0 + lane, pairs_total - 1);
            const long long b = p / per_b, rem = p - b * per_b;
            const int i = row_begin + (int)(rem / n), j = (int)(rem % n);
            float x[K + 1];
            pair_arguments<K>(__ldg(pts4 + b * n + i), __ldg(pts4 + b * n + j), refv + (b * n + i) * K, K, sigma_d,
                              factor_a, x);
#pragma unroll
            for (int s = 0; s < K; ++s) {
                clamped |= !(x[s] <= x_max_a);
                code[s] = tbl_encode(x[s], inv_ha, rows_a, 2);
            }
            x_d = x[K];
            code[K] = x_d <= x_max_d ? tbl_encode(x_d, inv_hd, rows_d, 4) : 0xffffffffu;  // beyond the table: exact
        }
        const int count = (int)min((long long)32, pairs_total - g0);
        // groups without a distance beyond the table (all of them, normally) run the branch-free loop
        if (!__any_sync(0xffffffffu, code[K] == 0xffffffffu)) {
            for (int e = 0; e < count; ++e) {
                __half2 best[C / 64], cur[C / 64];
#pragma unroll
                for (int s = 0; s < K; ++s) {  // geotransformer.py:66 max over the k angle slots
                    tbl_linear<C>(fa, __shfl_sync(0xffffffffu, code[s], e), lane, cur);
#pragma unroll
                    for (int q = 0; q < C / 64; ++q) best[q] = s == 0 ? cur[q] : __hmax2(best[q], cur[q]);
                }
                tbl_cubic<C>(fd, __shfl_sync(0xffffffffu, code[K], e), lane, cur);
                V o;
                __half2* op = reinterpret_cast<__half2*>(&o);
#pragma unroll
                for (int q = 0; q < C / 64; ++q) op[q] = __hadd2(cur[q], best[q]);  // :70
                tbl_store(reinterpret_cast<V*>(out + (size_t)(g0 + e) * C) + lane, o);
            }
        } else {
            for (int e = 0; e < count; ++e) {
                __half2 best[C / 64], cur[C / 64];
#pragma unroll
                for (int s = 0; s < K; ++s) {
                    tbl_linear<C>(fa, __shfl_sync(0xffffffffu, code[s], e), lane, cur);
#pragma unroll
                    for (int q = 0; q < C / 64; ++q) best[q] = s == 0 ? cur[q] : __hmax2(best[q], cur[q]);
                }
                const unsigned int code_d = __shfl_sync(0xffffffffu, code[K], e);
                const float xe = __shfl_sync(0xffffffffu, x_d, e);
                if (code_d != 0xffffffffu) tbl_cubic<C>(fd, code_d, lane, cur);
                else tbl_exact<C>(div_term, w_d, b_d, xe, lane, cur);
                V o;
                __half2* op = reinterpret_cast<__half2*>(&o);
#pragma unroll
                for (int q = 0; q < C / 64; ++q) op[q] = __hadd2(cur[q], best[q]);
                tbl_store(reinterpret_cast<V*>(out + (size_t)(g0 + e) * C) + lane, o);
            }
        }
    }
    if (clamped) *overflow = 1;
}

}  // namespace geotx

using namespace geotx;

extern "C" int64_t geotx_gse_table_bytes(int C, int rows_a, int rows_d) {
    if ((C != 256 && C != 128) || rows_a < 4 || rows_d < 4 || rows_a > 65535 || rows_d > 65535) return 0;
    return (int64_t)(rows_a + rows_d) * C * 2;
}

extern "C" int geotx_gse_build_tables(const float* div_term, const float* w_d, const float* b_d, const float* w_a,
                                      const float* b_a, int C, float h_a, float h_d, int rows_a, int rows_d,
                                      void* tables, void* stream) {
    GEOTX_REQUIRE(div_term && w_d && b_d && w_a && b_a && tables, "geotx_gse_build_tables: null pointer");
    GEOTX_REQUIRE((C == 256 || C == 128) && rows_a >= 4 && rows_d >= 4 && rows_a <= 65535 && rows_d <= 65535 &&
                      h_a > 0.f && h_d > 0.f, "geotx_gse_build_tables: C must be 128 or 256, 4..65535 rows per table");
    __half* t = (__half*)tables;
    gse_table_build_kernel<<<rows_a, C, 0, (cudaStream_t)stream>>>(div_term, w_a, b_a, C, h_a, 0, t);
    gse_table_build_kernel<<<rows_d, C, 0, (cudaStream_t)stream>>>(div_term, w_d, b_d, C, h_d, 1, t + (size_t)rows_a * C);
    return launch_status("geotx_gse_build_tables");
}

extern "C" int geotx_gse_embed_table(const void* geometry, int B, int n, int k, int row_begin, int row_count,
                                     const void* tables, int C, int rows_a, int rows_d, float h_a, float h_d,
                                     float sigma_d, float factor_a, const float* div_term, const float* w_d,
                                     const float* b_d, void* out_f16, int32_t* overflow, void* stream) {
    GEOTX_REQUIRE(geometry && tables && out_f16 && overflow && div_term && w_d && b_d, "geotx_gse_embed_table: null pointer");
    GEOTX_REQUIRE(B > 0 && n > 0 && k == 3 && (C == 256 || C == 128), "geotx_gse_embed_table: takes k = 3, C = 128 or 256");
    GEOTX_REQUIRE(row_begin >= 0 && row_count > 0 && row_begin + row_count <= n, "geotx_gse_embed_table: bad row slab");
    const size_t smem = (size_t)(rows_a + rows_d) * C * 2;
    GEOTX_REQUIRE(rows_a >= 4 && rows_d >= 4 && smem <= 220 * 1024, "geotx_gse_embed_table: tables of %zu B do not fit",
                  smem);
    const float4* pts4 = (const float4*)geometry;
    const float4* refv = pts4 + (size_t)B * n;
    const long long pairs = (long long)B * row_count * n;
    const int grid = sm_count() * (smem <= 110 * 1024 ? TBL_CTAS_PER_SM : 1);
    auto launch = [&](auto kernel) -> int {
        int rc = check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "geotx_gse_embed_table: shared memory");
        if (rc) return rc;
        kernel<<<grid, TBL_THREADS, smem, (cudaStream_t)stream>>>(pts4, refv, n, row_begin, row_count, pairs,
                                                                  (const __half*)tables, rows_a, rows_d, 1.0f / h_a,
                                                                  1.0f / h_d, sigma_d, factor_a, div_term, w_d, b_d,
                                                                  (__half*)out_f16, overflow);
        return 0;
    };
    int rc = C == 256 ? launch(gse_embed_table_kernel<3, 256>) : launch(gse_embed_table_kernel<3, 128>);
    if (rc) return rc;
    return launch_status("geotx_gse_embed_table");
}
