// K1: fused pairwise distance + per-row top-(k+1) selection (bit-exact k-NN indices), plus the
// elementwise d_idx / a_idx kernel behind GeometricStructureEmbedding.get_embedding_indices.
//
// Replaces /root/reference/geotransformer/modules/geotransformer/geotransformer.py:38-53 and
// /root/reference/geotransformer/modules/ops/pairwise_distance.py:4-31.  The N x N distance matrix is never
// stored: each warp owns one query row, streams the packed points (float4, coalesced 512 B per warp load)
// and keeps its k+1 best candidates in registers; the warp-level merge is k+1 rounds of shuffle arg-min.
#include "geotx_common.cuh"

namespace geotx {

// points (B*n, 3) fp32  ->  pts4 (B*n) float4 = (x, y, z, |p|^2 in the reference's rounding order)
// Also clears the word behind the geometry scratch in which knn_kernel leaves the largest squared distance.
__global__ void pack_points_kernel(const float* __restrict__ points, float4* __restrict__ pts4, int total,
                                   unsigned int* __restrict__ dmax2_bits) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0) *dmax2_bits = 0u;
    if (t >= total) return;
    float x = points[3 * t], y = points[3 * t + 1], z = points[3 * t + 2];
    pts4[t] = make_float4(x, y, z, sq_norm_ref(x, y, z));
}

constexpr int KNN_WARPS = 8;
constexpr int KK_MAX = GEOTX_MAX_K + 1;

__device__ __forceinline__ bool cand_less(float va, int ia, float vb, int ib) {
    return va < vb || (va == vb && ia < ib);
}

// One warp per query row.  Selection key is (clamped d2, index): sqrt is monotone, so this is the order the
// reference's topk over sqrt(d2) produces wherever that order is defined (tie-free rows, see synth.tie_free).
// KK = candidates kept per lane (>= k + 1): 4 for the shipped configurations (k = 3), KK_MAX otherwise.  The insertion
// network is unrolled over KK, so the small instance runs 80 instead of 154 instructions per scanned point and lane.
template <int KK>
__global__ void __launch_bounds__(KNN_WARPS * 32)
knn_kernel(const float4* __restrict__ pts4, int n, int k, int rows_total,
           long long* __restrict__ knn_out, float4* __restrict__ refv_out, unsigned int* __restrict__ dmax2_bits) {
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * KNN_WARPS + (threadIdx.x >> 5);
    if (row >= rows_total) return;
    const int b = row / n;
    const float4* __restrict__ base = pts4 + (size_t)b * n;
    const float4 pi = base[row - b * n];
    const int kk = k + 1;

    float bv[KK];
    int bi[KK];
#pragma unroll
    for (int q = 0; q < KK; ++q) { bv[q] = __int_as_float(0x7f800000); bi[q] = 0x7fffffff; }

    float dmax = 0.f;  // largest squared distance of the row: the range of the distance embedding (K2c)
    for (int j = lane; j < n; j += 32) {
        const float4 pj = __ldg(base + j);
        const float d = sqdist_ref(pi, pj);
        dmax = fmaxf(dmax, d);
        // strict '<' against the current worst kept candidate: j only grows within a lane, so equal values
        // keep the smaller index first
        bool worse_or_equal = true;
#pragma unroll
        for (int q = 0; q < KK; ++q)
            if (q == kk - 1) worse_or_equal = !(d < bv[q]);
        if (!worse_or_equal) {
            float cv = d;
            int ci = j;
#pragma unroll
            for (int q = 0; q < KK; ++q) {
                if (q < kk && cand_less(cv, ci, bv[q], bi[q])) {
                    float tv = bv[q]; int ti = bi[q];
                    bv[q] = cv; bi[q] = ci;
                    cv = tv; ci = ti;
                }
            }
        }
    }

#pragma unroll
    for (int off = 16; off > 0; off >>= 1) dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, off));
    if (lane == 0) atomicMax(dmax2_bits, __float_as_uint(dmax));  // d >= 0: the bit patterns order like the values

    // warp merge: pop the global minimum k+1 times; rank 0 is dropped (geotransformer.py:42 "[:, :, 1:]")
    for (int r = 0; r < kk; ++r) {
        float v = bv[0];
        int idx = bi[0];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, v, off);
            int oi = __shfl_xor_sync(0xffffffffu, idx, off);
            if (cand_less(ov, oi, v, idx)) { v = ov; idx = oi; }
        }
        if (bi[0] == idx) {  // indices are unique across lanes: exactly one lane owns the winner
#pragma unroll
            for (int q = 0; q < KK - 1; ++q) { bv[q] = bv[q + 1]; bi[q] = bi[q + 1]; }
            bv[KK - 1] = __int_as_float(0x7f800000);
            bi[KK - 1] = 0x7fffffff;
        }
        if (r >= 1 && lane == 0) {
            knn_out[(size_t)row * k + (r - 1)] = (long long)idx;
            const float4 pk = base[idx];
            // geotransformer.py:46  ref_vectors = knn_points - points
            refv_out[(size_t)row * k + (r - 1)] =
                make_float4(__fsub_rn(pk.x, pi.x), __fsub_rn(pk.y, pi.y), __fsub_rn(pk.z, pi.z), 0.f);
        }
    }
}

// d_idx (B,n,n) and a_idx (B,n,n,k): one thread per (i, j).  geotransformer.py:38-39, 46-53.
__global__ void __launch_bounds__(256)
indices_kernel(const float4* __restrict__ pts4, const float4* __restrict__ refv, int n, int k, long long pairs_total,
               float sigma_d, float factor_a, float* __restrict__ d_idx, float* __restrict__ a_idx) {
    long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pairs_total) return;
    long long bi_row = p / n;  // = b*n + i
    int j = (int)(p - bi_row * n);
    long long b = bi_row / n;
    const float4 pi = __ldg(pts4 + bi_row);
    const float4 pj = __ldg(pts4 + b * n + j);
    float x[GEOTX_MAX_K + 1];
    pair_arguments<GEOTX_MAX_K>(pi, pj, refv + bi_row * k, k, sigma_d, factor_a, x);
    d_idx[p] = x[GEOTX_MAX_K];
#pragma unroll
    for (int kk = 0; kk < GEOTX_MAX_K; ++kk)
        if (kk < k) a_idx[p * k + kk] = x[kk];
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
#include "host_util.h"
using namespace geotx;

extern "C" int64_t geotx_gse_geometry_bytes(int B, int n, int k) {
    if (B <= 0 || n <= 0 || k <= 0) return 0;
    return (int64_t)16 * B * n * (1 + (int64_t)k) + 16;  // + the largest squared distance (one word, 16-byte slot)
}

extern "C" int geotx_gse_knn(const float* points, int B, int n, int k, int64_t* knn_out, void* geometry,
                             void* stream) {
    GEOTX_REQUIRE(points && knn_out && geometry, "geotx_gse_knn: null pointer");
    GEOTX_REQUIRE(B > 0 && n > 0, "geotx_gse_knn: B and n must be positive (B=%d n=%d)", B, n);
    GEOTX_REQUIRE(k >= 1 && k <= GEOTX_MAX_ANGLE_K, "geotx_gse_knn: angle_k=%d outside [1,%d]", k, GEOTX_MAX_ANGLE_K);
    // torch.topk(k+1) raises for n < k+1 (geotransformer.py:42); so do we
    GEOTX_REQUIRE(n >= k + 1, "geotx_gse_knn: selected index k out of range (n=%d < k+1=%d)", n, k + 1);
    GEOTX_REQUIRE(((uintptr_t)geometry & 15) == 0, "geotx_gse_knn: geometry scratch must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    float4* pts4 = (float4*)geometry;
    float4* refv = pts4 + (size_t)B * n;
    int total = B * n;
    unsigned int* dmax2_bits = (unsigned int*)(refv + (size_t)total * k);
    pack_points_kernel<<<(total + 255) / 256, 256, 0, st>>>(points, pts4, total, dmax2_bits);
    if (k + 1 <= 4)
        knn_kernel<4><<<(total + KNN_WARPS - 1) / KNN_WARPS, KNN_WARPS * 32, 0, st>>>(
            pts4, n, k, total, (long long*)knn_out, refv, dmax2_bits);
    else
        knn_kernel<KK_MAX><<<(total + KNN_WARPS - 1) / KNN_WARPS, KNN_WARPS * 32, 0, st>>>(
            pts4, n, k, total, (long long*)knn_out, refv, dmax2_bits);
    return launch_status("geotx_gse_knn");
}

extern "C" int geotx_gse_indices(const void* geometry, int B, int n, int k, float sigma_d, float factor_a,
                                 float* d_idx, float* a_idx, void* stream) {
    GEOTX_REQUIRE(geometry && d_idx && a_idx, "geotx_gse_indices: null pointer");
    GEOTX_REQUIRE(B > 0 && n > 0 && k >= 1 && k <= GEOTX_MAX_ANGLE_K, "geotx_gse_indices: bad sizes");
    const float4* pts4 = (const float4*)geometry;
    const float4* refv = pts4 + (size_t)B * n;
    long long pairs = (long long)B * n * n;
    long long blocks = (pairs + 255) / 256;
    GEOTX_REQUIRE(blocks < 2147483647LL, "geotx_gse_indices: too many pairs");
    indices_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(pts4, refv, n, k, pairs, sigma_d, factor_a,
                                                                        d_idx, a_idx);
    return launch_status("geotx_gse_indices");
}
