// K9: radius neighbours in stack mode: for every query point the support points of the same batch
// element closer than `radius`, nearest first, truncated to `limit`, padded with the sentinel (total support count).
//
// Replaces the reference's CPU collate operator (geotransformer/extensions/cpu/radius_neighbors/radius_neighbors_cpu.cpp:
// 3-104: one nanoflann KD-tree per batch element, radiusSearch with sorted results) followed by the truncation of
// modules/ops/radius_search.py:24-27.  It runs 3 times per pyramid stage in utils/data.py:33-72, on the host, inside
// the data loader.  On a B200 no tree is needed: the exact brute-force scan of a 15k x 15k stage is 0.2 G distance
// evaluations, ~100 us.
//
// One warp per query.  The CTA stages the support points of the batch element through shared memory in tiles; a lane
// tests one support point per step with the reference's own arithmetic (nanoflann L2_Simple_Adaptor::evalMetric:
// d = ((qx-sx)^2 + (qy-sy)^2) + (qz-sz)^2 in float, in-radius iff d < radius*radius, nanoflann.hpp:249-251,432-439),
// hits are compacted by ballot into the warp's candidate buffer as 64-bit keys (distance bits, support index); when
// the buffer fills up it is bitonic-sorted and cut back to `limit`.  Keys are unique, so the result is deterministic;
// equal distances are ordered by index (std::sort in the reference leaves ties unspecified).
#include "geotx_common.cuh"

namespace geotx {

constexpr int RN_WARPS = 8;
constexpr int RN_TILE = 1024;  // support points per shared-memory tile (16 KB)

// Ascending bitonic sort of buf[0, len) (len a power of two >= 64) by one warp.
__device__ __forceinline__ void warp_bitonic_sort(unsigned long long* buf, int len, int lane) {
    for (int k = 2; k <= len; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = lane; t < (len >> 1); t += 32) {
                const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int hi = lo | j;
                const bool up = (lo & k) == 0;
                const unsigned long long a = buf[lo], b = buf[hi];
                if ((a > b) == up) { buf[lo] = b; buf[hi] = a; }
            }
            __syncwarp();
        }
    }
}

__global__ void __launch_bounds__(RN_WARPS * 32)
radius_neighbors_kernel(const float* __restrict__ q_points, int q_begin, int q_count, const float* __restrict__ s_points,
                        int s_begin, int s_count, long long sentinel, float r2, int limit, int buf_len,
                        long long* __restrict__ neighbors, int* __restrict__ counts, int* __restrict__ max_count) {
    extern __shared__ __align__(16) unsigned char rn_smem[];
    float4* tile = reinterpret_cast<float4*>(rn_smem);
    unsigned long long* bufs = reinterpret_cast<unsigned long long*>(rn_smem + RN_TILE * sizeof(float4));
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long* buf = bufs + (size_t)wid * buf_len;
    const int q = blockIdx.x * RN_WARPS + wid;  // query index inside the batch element
    const bool live = q < q_count;
    float qx = 0.f, qy = 0.f, qz = 0.f;
    if (live) {
        const float* p = q_points + 3 * (size_t)(q_begin + q);
        qx = p[0]; qy = p[1]; qz = p[2];
    }
    int filled = 0;  // keys in buf (warp-uniform)
    int total = 0;   // in-radius support points seen so far
    for (int t0 = 0; t0 < s_count; t0 += RN_TILE) {
        const int cnt = min(RN_TILE, s_count - t0);
        __syncthreads();
        for (int s = threadIdx.x; s < cnt; s += RN_WARPS * 32) {
            const float* p = s_points + 3 * (size_t)(s_begin + t0 + s);
            tile[s] = make_float4(p[0], p[1], p[2], 0.f);
        }
        __syncthreads();
        if (!live) continue;
        for (int s = lane; s < ((cnt + 31) & ~31); s += 32) {
            bool hit = false;
            float d = 0.f;
            if (s < cnt) {
                const float4 sp = tile[s];
                const float dx = __fsub_rn(qx, sp.x), dy = __fsub_rn(qy, sp.y), dz = __fsub_rn(qz, sp.z);
                d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                hit = d < r2;
            }
            const unsigned int ballot = __ballot_sync(0xffffffffu, hit);
            if (ballot == 0u) continue;
            const int n_hit = __popc(ballot);
            if (filled + n_hit > buf_len) {  // make room: keep the best `limit`
                for (int t = filled + lane; t < buf_len; t += 32) buf[t] = ~0ull;
                __syncwarp();
                warp_bitonic_sort(buf, buf_len, lane);
                filled = min(filled, limit);
            }
            if (hit)
                buf[filled + __popc(ballot & ((1u << lane) - 1u))] =
                    ((unsigned long long)__float_as_uint(d) << 32) | (unsigned int)(t0 + s);  // d >= 0: bits are ordered
            filled += n_hit;
            total += n_hit;
            __syncwarp();
        }
    }
    if (!live) return;
    for (int t = filled + lane; t < buf_len; t += 32) buf[t] = ~0ull;
    __syncwarp();
    warp_bitonic_sort(buf, buf_len, lane);
    const int valid = min(min(filled, limit), total);
    long long* out = neighbors + (size_t)(q_begin + q) * limit;
    for (int t = lane; t < limit; t += 32)
        out[t] = t < valid ? (long long)(s_begin + (int)(buf[t] & 0xffffffffull)) : sentinel;
    if (lane == 0) {
        counts[q_begin + q] = total;
        atomicMax(max_count, total);
    }
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
#include "host_util.h"
using namespace geotx;

extern "C" int geotx_radius_neighbors(const float* q_points, const float* s_points, const int64_t* host_q_lengths,
                                      const int64_t* host_s_lengths, int batch, float radius, int limit,
                                      int64_t* neighbors, int32_t* counts, int32_t* max_count, void* stream) {
    GEOTX_REQUIRE(q_points && s_points && host_q_lengths && host_s_lengths && neighbors && counts && max_count,
                  "geotx_radius_neighbors: null pointer");
    GEOTX_REQUIRE(batch > 0 && limit > 0 && limit <= GEOTX_MAX_NEIGHBOR_LIMIT,
                  "geotx_radius_neighbors: bad batch=%d or limit=%d (max %d)", batch, limit, GEOTX_MAX_NEIGHBOR_LIMIT);
    int64_t nq = 0, ns = 0;
    for (int b = 0; b < batch; ++b) {
        GEOTX_REQUIRE(host_q_lengths[b] >= 0 && host_s_lengths[b] >= 0, "geotx_radius_neighbors: negative length");
        nq += host_q_lengths[b];
        ns += host_s_lengths[b];
    }
    GEOTX_REQUIRE(nq < 2147483647LL && ns < 2147483647LL, "geotx_radius_neighbors: too many points");
    cudaStream_t st = (cudaStream_t)stream;
    int rc = check_cuda(cudaMemsetAsync(max_count, 0, sizeof(int32_t), st), "geotx_radius_neighbors: memset");
    if (rc) return rc;
    int buf_len = 64;
    while (buf_len < 2 * limit) buf_len <<= 1;
    const size_t smem = RN_TILE * sizeof(float4) + (size_t)RN_WARPS * buf_len * sizeof(unsigned long long);
    if (smem > 48 * 1024) {
        rc = check_cuda(cudaFuncSetAttribute(radius_neighbors_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem), "geotx_radius_neighbors: shared memory");
        if (rc) return rc;
    }
    const float r2 = radius * radius;  // radius_neighbors_cpu.cpp:12
    int64_t q0 = 0, s0 = 0;
    for (int b = 0; b < batch; ++b) {
        const int qc = (int)host_q_lengths[b], sc = (int)host_s_lengths[b];
        if (qc > 0)
            radius_neighbors_kernel<<<(qc + RN_WARPS - 1) / RN_WARPS, RN_WARPS * 32, smem, st>>>(
                q_points, (int)q0, qc, s_points, (int)s0, sc, (long long)ns, r2, limit, buf_len, (long long*)neighbors,
                counts, max_count);
        q0 += qc;
        s0 += sc;
    }
    return launch_status("geotx_radius_neighbors");
}
