// Shared device helpers for the GeoTransformer hot-path kernels (sm_100a only).
//
// Numerics contract: the squared distance must reproduce the reference's torch-CPU
// fp32 op sequence bit for bit, because the k-NN indices derived from it are compared bit-exactly:
//     xy = fma(x2, y2, fma(x1, y1, x0*y0))        /root/reference/.../ops/pairwise_distance.py:23 (MKL sgemm K=3)
//     sq = (x0*x0 + x1*x1) + x2*x2                 pairwise_distance.py:27-28 (no fma)
//     d2 = max(0, (sq_i - 2*xy) + sq_j)            pairwise_distance.py:29-30 (left to right)
// All of it is written with explicit round-to-nearest intrinsics so that nvcc can never contract it.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#define GEOTX_MAX_K 7  // angle_k <= 7 (k+1 <= 8 candidates kept per row)

namespace geotx {

__device__ __forceinline__ float sq_norm_ref(float x, float y, float z) {
    return __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z));
}

// p = (x, y, z, sq_norm_ref(x,y,z))
__device__ __forceinline__ float sqdist_ref(const float4& pi, const float4& pj) {
    float xy = __fmaf_rn(pi.z, pj.z, __fmaf_rn(pi.y, pj.y, __fmul_rn(pi.x, pj.x)));
    float d2 = __fadd_rn(__fsub_rn(pi.w, __fmul_rn(2.0f, xy)), pj.w);
    return fmaxf(d2, 0.0f);
}

// geotransformer.py:47-52: angle between ref = p_knn - p_i and anc = p_j - p_i, atan2(|ref x anc|, ref . anc)
__device__ __forceinline__ float triplet_angle(float rx, float ry, float rz, float ax, float ay, float az) {
    float cx = __fsub_rn(__fmul_rn(ry, az), __fmul_rn(rz, ay));
    float cy = __fsub_rn(__fmul_rn(rz, ax), __fmul_rn(rx, az));
    float cz = __fsub_rn(__fmul_rn(rx, ay), __fmul_rn(ry, ax));
    float s = sqrtf(cx * cx + cy * cy + cz * cz);
    // torch.sum starts from +0, so an all-zero dot product (j == i: anc = 0) is +0 even when the products are -0;
    // atan2(0, -0) would be pi instead of the reference's 0.  "+ 0.0f" maps -0 to +0 and changes nothing else.
    float c = __fadd_rn(rx * ax + ry * ay + rz * az, 0.0f);
    return atan2f(s, c);
}

// The four embedding arguments of one (i, j) pair: x[0..k-1] = angle_kk * factor_a, x[k] = dist / sigma_d.
// refv: k float4 per point (p_knn - p_i, .w unused), written by the k-NN kernel.
template <int MAXK>
__device__ __forceinline__ void pair_arguments(const float4& pi, const float4& pj, const float4* __restrict__ refv_i,
                                               int k, float sigma_d, float factor_a, float* x /*MAXK+1*/) {
    float ax = __fsub_rn(pj.x, pi.x), ay = __fsub_rn(pj.y, pi.y), az = __fsub_rn(pj.z, pi.z);
#pragma unroll
    for (int kk = 0; kk < MAXK; ++kk) {
        if (kk < k) {
            float4 r = __ldg(refv_i + kk);
            x[kk] = __fmul_rn(triplet_angle(r.x, r.y, r.z, ax, ay, az), factor_a);
        }
    }
    x[MAXK] = __fdiv_rn(__fsqrt_rn(sqdist_ref(pi, pj)), sigma_d);
}

// One of the k+1 embedding arguments of a pair: s < k -> angle s * factor_a, s == k -> dist / sigma_d (same arithmetic
// as pair_arguments; lets cooperating threads split the arguments of a pair between them).
__device__ __forceinline__ float pair_argument_one(const float4& pi, const float4& pj, const float4* __restrict__ refv_i,
                                                   int k, int s, float sigma_d, float factor_a) {
    if (s < k) {
        const float ax = __fsub_rn(pj.x, pi.x), ay = __fsub_rn(pj.y, pi.y), az = __fsub_rn(pj.z, pi.z);
        const float4 r = __ldg(refv_i + s);
        return __fmul_rn(triplet_angle(r.x, r.y, r.z, ax, ay, az), factor_a);
    }
    return __fdiv_rn(__fsqrt_rn(sqdist_ref(pi, pj)), sigma_d);
}

// ---- small PTX wrappers -------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// sin(y), cos(y) on the FMA pipe (no MUFU): y = k*pi + r with k = rint(y/pi) by the 1.5*2^23 trick, |r| <= pi/2,
// degree-7 odd / degree-6 even least-squares polynomials on Chebyshev nodes (max error 1e-6 / 8e-6 on the reduced
// range), sign (-1)^k applied through the sign bit.  End-to-end |error| <= 1.1e-5 for |y| <= 60 (tools/ fit script in
// DESIGN.md), i.e. 200x below the bf16 rounding the result goes through.  13 FMA-pipe + 3 ALU instructions per pair;
// the fused embedding kernel uses it for half of the frequencies because MUFU (16/clk/SM) is its co-bottleneck.
__device__ __forceinline__ void sincos_fma(float y, float& s, float& c) {
    const float t = __fmaf_rn(y, 0.31830987f, 12582912.0f);
    const float k = t - 12582912.0f;
    const float r = __fmaf_rn(k, -3.14159274f, y);
    const uint32_t sign = __float_as_uint(t) << 31;  // parity of k
    const float r2 = r * r;
    float sp = __fmaf_rn(r2, -1.848342508e-4f, 8.312028833e-3f);
    sp = __fmaf_rn(r2, sp, -0.1666564941f);
    const float sv = __fmaf_rn(r * r2, sp, r);
    float cp = __fmaf_rn(r2, -1.275219372e-3f, 4.150512442e-2f);
    cp = __fmaf_rn(r2, cp, -0.4999339879f);
    const float cv = __fmaf_rn(r2, cp, 1.0f);
    s = __uint_as_float(__float_as_uint(sv) ^ sign);
    c = __uint_as_float(__float_as_uint(cv) ^ sign);
}

// 32-byte store (STG.256, sm_100): a full sector per thread, so row-strided writers do not emit half sectors.
// L1::no_allocate (STG.E.NA): the lines are written once and never read by this SM; without the hint the embedding
// kernel is 6 % slower (0.468 -> 0.439 ms at n = 1024; .cs / .cg make no difference) -- the stores share the LSU/L1
// path with the producers' shared-memory traffic.  -DK2_ST_HINT=<qualifier> overrides it for timing experiments.
#ifndef K2_ST_HINT
#define K2_ST_HINT .L1::no_allocate
#endif
#define GEOTX_STR2(x) #x
#define GEOTX_STR(x) GEOTX_STR2(x)
__device__ __forceinline__ void st_global_v8(void* p, const uint32_t* v) {
    asm volatile("st.global" GEOTX_STR(K2_ST_HINT) ".v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(v[0]),
                 "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}

}  // namespace geotx
