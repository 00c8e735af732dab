// K9: radius neighbours in stack mode: for every query point the support points of the same batch
// element closer than `radius`, nearest first, truncated to `limit`, padded with the sentinel (total support count).
//
// Replaces the reference's CPU collate operator (geotransformer/extensions/cpu/radius_neighbors/radius_neighbors_cpu.cpp:
// 3-104: one nanoflann KD-tree per batch element, radiusSearch with sorted results) followed by the truncation of
// modules/ops/radius_search.py:24-27.  It runs 3 times per pyramid stage in utils/data.py:33-72, on the host, inside
// the data loader.  On a B200 no tree is needed: the exact brute-force scan of a 15k x 15k stage is 0.2 G distance
// evaluations, ~100 us.
//
// One warp per 4 queries.  The CTA stages the support points of the batch element through shared memory in tiles; a
// lane reads one support point per step and tests it against the warp's 4 queries with the reference's own arithmetic
// (nanoflann L2_Simple_Adaptor::evalMetric: d = ((qx-sx)^2 + (qy-sy)^2) + (qz-sz)^2 in float, in-radius iff
// d < radius*radius, nanoflann.hpp:249-251,432-439); hits are compacted by ballot into the query's candidate buffer
// as 64-bit keys (distance bits, support index); when a buffer fills up it is bitonic-sorted and cut back to `limit`.  Keys are unique, so the result is deterministic;
// equal distances are ordered by index (std::sort in the reference leaves ties unspecified).
#include "geotx_common.cuh"

namespace geotx {

constexpr int RN_WARPS = 4;
constexpr int RN_Q = 4;        // queries per warp: one shared-memory read of a support point serves RN_Q distance tests
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

// Appends the lanes' hits of one step to a query's candidate buffer (all lanes of the warp call it together).
__device__ __forceinline__ void rn_append(unsigned long long* buf, int buf_len, int limit, int lane, unsigned int ballot,
                                          bool hit, float d, int support, int& filled, int& total) {
    const int n_hit = __popc(ballot);
    if (filled + n_hit > buf_len) {  // make room: keep the best `limit`
        for (int t = filled + lane; t < buf_len; t += 32) buf[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(buf, buf_len, lane);
        filled = min(filled, limit);
    }
    if (hit)
        buf[filled + __popc(ballot & ((1u << lane) - 1u))] =
            ((unsigned long long)__float_as_uint(d) << 32) | (unsigned int)support;  // d >= 0: bits are ordered
    filled += n_hit;
    total += n_hit;
    __syncwarp();
}

__global__ void __launch_bounds__(RN_WARPS * 32)
radius_neighbors_kernel(const float* __restrict__ q_points, int q_begin, int q_count, const float* __restrict__ s_points,
                        int s_begin, int s_count, long long sentinel, float r2, int limit, int buf_len,
                        long long* __restrict__ neighbors, int* __restrict__ counts, int* __restrict__ max_count) {
    extern __shared__ __align__(16) unsigned char rn_smem[];
    float4* tile = reinterpret_cast<float4*>(rn_smem);
    unsigned long long* bufs = reinterpret_cast<unsigned long long*>(rn_smem + RN_TILE * sizeof(float4));
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long* buf = bufs + (size_t)wid * RN_Q * buf_len;  // RN_Q buffers of buf_len keys
    const int q0 = (blockIdx.x * RN_WARPS + wid) * RN_Q;            // first query of this warp inside the batch element
    float qx[RN_Q], qy[RN_Q], qz[RN_Q];
    int filled[RN_Q], total[RN_Q];
#pragma unroll
    for (int k = 0; k < RN_Q; ++k) {
        const int q = min(q0 + k, q_count - 1);  // surplus slots repeat the last query and are not written back
        const float* p = q_points + 3 * (size_t)(q_begin + q);
        qx[k] = p[0]; qy[k] = p[1]; qz[k] = p[2];
        filled[k] = 0;
        total[k] = 0;
    }
    const bool live = q0 < q_count;
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
            const bool in_tile = s < cnt;
            const float4 sp = tile[in_tile ? s : 0];
            float d[RN_Q];
            unsigned int ballot[RN_Q], any = 0u;
#pragma unroll
            for (int k = 0; k < RN_Q; ++k) {
                const float dx = __fsub_rn(qx[k], sp.x), dy = __fsub_rn(qy[k], sp.y), dz = __fsub_rn(qz[k], sp.z);
                d[k] = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                ballot[k] = __ballot_sync(0xffffffffu, in_tile && d[k] < r2);
                any |= ballot[k];
            }
            if (any == 0u) continue;
#pragma unroll
            for (int k = 0; k < RN_Q; ++k)
                if (ballot[k])
                    rn_append(buf + (size_t)k * buf_len, buf_len, limit, lane, ballot[k], (ballot[k] >> lane) & 1u, d[k],
                              t0 + s, filled[k], total[k]);
        }
    }
    if (!live) return;
#pragma unroll
    for (int k = 0; k < RN_Q; ++k) {
        if (q0 + k >= q_count) break;
        unsigned long long* b = buf + (size_t)k * buf_len;
        for (int t = filled[k] + lane; t < buf_len; t += 32) b[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(b, buf_len, lane);
        const int valid = min(filled[k], limit);
        long long* out = neighbors + (size_t)(q_begin + q0 + k) * limit;
        for (int t = lane; t < limit; t += 32)
            out[t] = t < valid ? (long long)(s_begin + (int)(b[t] & 0xffffffffull)) : sentinel;
        if (lane == 0) {
            counts[q_begin + q0 + k] = total[k];
            atomicMax(max_count, total[k]);
        }
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
    const size_t smem = RN_TILE * sizeof(float4) + (size_t)RN_WARPS * RN_Q * buf_len * sizeof(unsigned long long);
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
            radius_neighbors_kernel<<<(qc + RN_WARPS * RN_Q - 1) / (RN_WARPS * RN_Q), RN_WARPS * 32, smem, st>>>(
                q_points, (int)q0, qc, s_points, (int)s0, sc, (long long)ns, r2, limit, buf_len, (long long*)neighbors,
                counts, max_count);
        q0 += qc;
        s0 += sc;
    }
    return launch_status("geotx_radius_neighbors");
}
