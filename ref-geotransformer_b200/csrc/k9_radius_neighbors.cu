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

// Ascending bitonic sort of buf[0, len) (len a power of two >= 2) by one warp.
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
        int sort_len = 32;
        while (sort_len < filled[k]) sort_len <<= 1;
        for (int t = filled[k] + lane; t < sort_len; t += 32) b[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(b, sort_len, lane);
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

// ---- cell-list path ---------------------------------------------------------------------------------------------
// The brute-force scan above tests every support point of the batch element; at the pyramid's sizes (20 000 points in a
// 3 m scene, radius 0.0625 m) only ~1 in 10 000 tests hits.  The cell-list path bins the support points of a batch
// element into a uniform grid of cells slightly larger than the radius (per-cell linked lists: head[cell] and a link
// stored next to every point's coordinates, built with one atomicExch per point -- no sort, no scan) and a query only visits its 27 surrounding cells: lane c of
// the query's warp walks the list of cell c.  Candidates go through the SAME distance arithmetic, hit test, 64-bit
// (distance, index) keys and bitonic select as the brute-force path, so the results are identical (the order in which
// candidates are met does not matter: keys are unique and sorted).  A cell edge of 1.001 x radius (or larger, when the
// grid would not fit `max_cells`) covers the rounding of the cell index computation.
constexpr int RNG_SLOTS = 8;  // batch elements searched by one set of launches (blockIdx.y)

struct RnBatch {  // by value in the kernel parameters: the lengths live on the host, as in the reference
    int n;
    int q_begin[RNG_SLOTS], q_count[RNG_SLOTS], s_begin[RNG_SLOTS], s_count[RNG_SLOTS];
};

struct RnGrid {  // one grid per slot
    unsigned int* bounds;  // [2][RNG_SLOTS][3] order-preserving encodings: min x,y,z of every slot, then max x,y,z
    float* params;         // [RNG_SLOTS][8] origin x,y,z, 1/cell, then nx, ny, nz as int bits
    int* head;             // [RNG_SLOTS][max_cells] first point of the cell's list (index inside the element), -1: empty
    long long max_cells;
    float4* cell_points;   // [stacked support points] x, y, z and (as int bits) the next point of the cell's list
};

__device__ __forceinline__ unsigned int rn_encode(float f) {
    const unsigned int b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float rn_decode(unsigned int e) {
    return __uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e);
}

__global__ void rn_bounds_kernel(const float* __restrict__ s_points, RnBatch b, RnGrid g) {
    const int e = blockIdx.y, s_begin = b.s_begin[e], s_count = b.s_count[e];
    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < s_count; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const float v = s_points[3 * (size_t)(s_begin + i) + d];
            lo[d] = fminf(lo[d], v);
            hi[d] = fmaxf(hi[d], v);
        }
    }
#pragma unroll
    for (int d = 0; d < 3; ++d) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            lo[d] = fminf(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], off));
            hi[d] = fmaxf(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], off));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&g.bounds[3 * e + d], rn_encode(lo[d]));
            atomicMax(&g.bounds[3 * (RNG_SLOTS + e) + d], rn_encode(hi[d]));
        }
    }
}

// One thread per slot: the grid of its batch element.
__global__ void rn_params_kernel(float radius, RnBatch b, RnGrid g) {
    const int e = threadIdx.x;
    if (e >= b.n) return;
    float* params = g.params + 8 * e;
    float lo[3] = {0.f, 0.f, 0.f}, ext[3] = {0.f, 0.f, 0.f};
    if (b.s_count[e] > 0) {
        for (int d = 0; d < 3; ++d) {
            lo[d] = rn_decode(g.bounds[3 * e + d]);
            ext[d] = fmaxf(rn_decode(g.bounds[3 * (RNG_SLOTS + e) + d]) - lo[d], 0.f);
        }
    }
    float cell = fmaxf(radius * 1.001f, 1e-30f);
    int n[3];
    for (;;) {
        long long total = 1;
        bool fits = true;
        for (int d = 0; d < 3; ++d) {
            const float cnt = floorf(ext[d] / cell) + 1.f;
            if (!(cnt < 2048.f)) { fits = false; break; }  // keeps the index rounding far below the 0.001 margin
            n[d] = (int)cnt;
            total *= n[d];
        }
        if (fits && total <= g.max_cells) break;
        cell *= 1.25f;
    }
    for (int d = 0; d < 3; ++d) {
        params[d] = lo[d];
        params[4 + d] = __int_as_float(n[d]);
    }
    params[3] = 1.0f / cell;
}

__device__ __forceinline__ void rn_cell_of(const float* __restrict__ params, float x, float y, float z, int (&c)[3]) {
    c[0] = (int)floorf((x - params[0]) * params[3]);
    c[1] = (int)floorf((y - params[1]) * params[3]);
    c[2] = (int)floorf((z - params[2]) * params[3]);
}

__global__ void rn_build_kernel(const float* __restrict__ s_points, RnBatch b, RnGrid g) {
    const int e = blockIdx.y, s_begin = b.s_begin[e], s_count = b.s_count[e];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s_count) return;
    const float* params = g.params + 8 * e;
    const float* p = s_points + 3 * (size_t)(s_begin + i);
    int c[3];
    rn_cell_of(params, p[0], p[1], p[2], c);
    const int nx = __float_as_int(params[4]), ny = __float_as_int(params[5]), nz = __float_as_int(params[6]);
    c[0] = min(max(c[0], 0), nx - 1);  // support points are inside the box; the clamp only absorbs rounding at the faces
    c[1] = min(max(c[1], 0), ny - 1);
    c[2] = min(max(c[2], 0), nz - 1);
    const int next = atomicExch(&g.head[(size_t)e * g.max_cells + ((size_t)c[2] * ny + c[1]) * nx + c[0]], i);
    g.cell_points[s_begin + i] = make_float4(p[0], p[1], p[2], __int_as_float(next));  // one 16-byte load per hop
}

constexpr int RNG_WARPS = 8;  // queries per CTA, one warp each

__global__ void __launch_bounds__(RNG_WARPS * 32)
rn_grid_query_kernel(const float* __restrict__ q_points, RnBatch b, long long sentinel, float r2, int limit, int buf_len,
                     RnGrid g, long long* __restrict__ neighbors, int* __restrict__ counts,
                     int* __restrict__ max_count) {
    const int e = blockIdx.y, q_begin = b.q_begin[e], q_count = b.q_count[e], s_begin = b.s_begin[e];
    const float* params = g.params + 8 * e;
    const int* head = g.head + (size_t)e * g.max_cells;
    const float4* cell_points = g.cell_points + s_begin;
    extern __shared__ __align__(16) unsigned char rng_smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long* buf = reinterpret_cast<unsigned long long*>(rng_smem) + (size_t)wid * buf_len;
    const int q = blockIdx.x * RNG_WARPS + wid;
    if (q >= q_count) return;
    const float* qp = q_points + 3 * (size_t)(q_begin + q);
    const float qx = qp[0], qy = qp[1], qz = qp[2];
    const int nx = __float_as_int(params[4]), ny = __float_as_int(params[5]), nz = __float_as_int(params[6]);
    int c[3];
    rn_cell_of(params, qx, qy, qz, c);
    // A query may lie outside the support box (the two clouds of a search differ): its own cell is not clamped, cells
    // outside the grid are empty.  A support point clamped onto a face cell is still within one cell of where the
    // unclamped index would put it.
    int cur = -1;
    if (lane < 27) {
        const int cx = c[0] + lane % 3 - 1, cy = c[1] + (lane / 3) % 3 - 1, cz = c[2] + lane / 9 - 1;
        if (cx >= 0 && cx < nx && cy >= 0 && cy < ny && cz >= 0 && cz < nz)
            cur = head[((size_t)cz * ny + cy) * nx + cx];
    }
    int filled = 0, total = 0;
    while (__any_sync(0xffffffffu, cur >= 0)) {
        float d = 0.f;
        bool hit = false;
        const int cand = cur;
        if (cand >= 0) {
            const float4 sp = cell_points[cand];
            const float dx = __fsub_rn(qx, sp.x), dy = __fsub_rn(qy, sp.y), dz = __fsub_rn(qz, sp.z);
            d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            hit = d < r2;
            cur = __float_as_int(sp.w);
        }
        const unsigned int ballot = __ballot_sync(0xffffffffu, hit);
        if (ballot) rn_append(buf, buf_len, limit, lane, ballot, hit, d, cand, filled, total);
    }
    int sort_len = 32;  // a query of the pyramid keeps ~10 hits: sort what there is, not the whole buffer
    while (sort_len < filled) sort_len <<= 1;
    for (int t = filled + lane; t < sort_len; t += 32) buf[t] = ~0ull;
    __syncwarp();
    warp_bitonic_sort(buf, sort_len, lane);
    const int valid = min(filled, limit);
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

static size_t rn_align(size_t x) { return (x + 255) & ~(size_t)255; }
static size_t rn_grid_layout(int64_t max_support, int64_t max_cells, char* base, RnGrid* g) {
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += rn_align(bytes); return p; };
    char* a = take(2 * RNG_SLOTS * 3 * sizeof(unsigned int));
    char* b = take(RNG_SLOTS * 8 * sizeof(float));
    char* c = take((size_t)RNG_SLOTS * max_cells * sizeof(int));
    char* d = take((size_t)max_support * sizeof(float4));
    if (g) {
        g->bounds = (unsigned int*)a;
        g->params = (float*)b;
        g->head = (int*)c;
        g->max_cells = max_cells;
        g->cell_points = (float4*)d;
    }
    return off;
}

extern "C" int64_t geotx_radius_neighbors_workspace_bytes(int64_t max_support_points, int64_t max_cells) {
    if (max_support_points < 0 || max_cells < 1 || max_cells > 268435456LL || max_support_points > 2147483647LL) return 0;
    return (int64_t)rn_grid_layout(max_support_points, max_cells, nullptr, nullptr);
}

extern "C" int geotx_radius_neighbors_grid(const float* q_points, const float* s_points, const int64_t* host_q_lengths,
                                           const int64_t* host_s_lengths, int batch, float radius, int limit,
                                           int64_t* neighbors, int32_t* counts, int32_t* max_count, void* workspace,
                                           int64_t max_support_points, int64_t max_cells, void* stream) {
    GEOTX_REQUIRE(q_points && s_points && host_q_lengths && host_s_lengths && neighbors && counts && max_count && workspace,
                  "geotx_radius_neighbors_grid: null pointer");
    GEOTX_REQUIRE(batch > 0 && limit > 0 && limit <= GEOTX_MAX_NEIGHBOR_LIMIT,
                  "geotx_radius_neighbors_grid: bad batch=%d or limit=%d (max %d)", batch, limit, GEOTX_MAX_NEIGHBOR_LIMIT);
    GEOTX_REQUIRE(max_cells >= 1 && max_cells <= 268435456LL && radius > 0.f,
                  "geotx_radius_neighbors_grid: bad max_cells=%lld or radius", (long long)max_cells);
    int64_t nq = 0, ns = 0;
    for (int b = 0; b < batch; ++b) {
        GEOTX_REQUIRE(host_q_lengths[b] >= 0 && host_s_lengths[b] >= 0, "geotx_radius_neighbors_grid: negative length");
        nq += host_q_lengths[b];
        ns += host_s_lengths[b];
    }
    GEOTX_REQUIRE(nq < 2147483647LL && ns < 2147483647LL && ns <= max_support_points,
                  "geotx_radius_neighbors_grid: %lld support points, workspace sized for %lld", (long long)ns,
                  (long long)max_support_points);
    cudaStream_t st = (cudaStream_t)stream;
    RnGrid g;
    rn_grid_layout(max_support_points, max_cells, (char*)workspace, &g);
    int rc = check_cuda(cudaMemsetAsync(max_count, 0, sizeof(int32_t), st), "geotx_radius_neighbors_grid: memset");
    if (rc) return rc;
    int buf_len = 64;
    while (buf_len < 2 * limit) buf_len <<= 1;
    const size_t smem = (size_t)RNG_WARPS * buf_len * sizeof(unsigned long long);
    if (smem > 48 * 1024) {
        rc = check_cuda(cudaFuncSetAttribute(rn_grid_query_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                        "geotx_radius_neighbors_grid: shared memory");
        if (rc) return rc;
    }
    const float r2 = radius * radius;  // radius_neighbors_cpu.cpp:12
    int64_t q0 = 0, s0 = 0;
    for (int first = 0; first < batch; first += RNG_SLOTS) {  // RNG_SLOTS batch elements per set of launches
        RnBatch b;
        b.n = batch - first < RNG_SLOTS ? batch - first : RNG_SLOTS;
        int q_max = 0, s_max = 0;
        for (int e = 0; e < RNG_SLOTS; ++e) {
            const bool live = e < b.n;
            b.q_begin[e] = (int)q0;
            b.s_begin[e] = (int)s0;
            b.q_count[e] = live ? (int)host_q_lengths[first + e] : 0;
            b.s_count[e] = live ? (int)host_s_lengths[first + e] : 0;
            q0 += b.q_count[e];
            s0 += b.s_count[e];
            q_max = b.q_count[e] > q_max ? b.q_count[e] : q_max;
            s_max = b.s_count[e] > s_max ? b.s_count[e] : s_max;
        }
        if (q_max == 0) continue;
        rc = check_cuda(cudaMemsetAsync(g.bounds, 0xff, RNG_SLOTS * 3 * sizeof(unsigned int), st), "memset bounds");
        if (!rc) rc = check_cuda(cudaMemsetAsync(g.bounds + RNG_SLOTS * 3, 0, RNG_SLOTS * 3 * sizeof(unsigned int), st),
                                 "memset bounds");
        if (!rc) rc = check_cuda(cudaMemsetAsync(g.head, 0xff, (size_t)b.n * max_cells * sizeof(int), st), "memset head");
        if (rc) return rc;
        if (s_max > 0) {
            const int blocks = (s_max + 255) / 256;
            rn_bounds_kernel<<<dim3(blocks < sm_count() ? blocks : sm_count(), b.n), 256, 0, st>>>(s_points, b, g);
        }
        rn_params_kernel<<<1, 32, 0, st>>>(radius, b, g);
        if (s_max > 0) rn_build_kernel<<<dim3((s_max + 255) / 256, b.n), 256, 0, st>>>(s_points, b, g);
        rn_grid_query_kernel<<<dim3((q_max + RNG_WARPS - 1) / RNG_WARPS, b.n), RNG_WARPS * 32, smem, st>>>(
            q_points, b, (long long)ns, r2, limit, buf_len, g, (long long*)neighbors, counts, max_count);
    }
    return launch_status("geotx_radius_neighbors_grid");
}
