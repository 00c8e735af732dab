// K10: grid (voxel barycentre) subsampling in stack mode.
//
// Replaces the reference's CPU collate operator geotransformer/extensions/cpu/grid_subsampling/grid_subsampling_cpu.cpp:
// 3-79 (exported as ext.grid_subsampling; modules/ops/grid_subsample.py:9-22; called once per pyramid stage from
// utils/data.py:22-30): per batch element, points are binned into voxels of size `voxel_size` anchored at
// floor(min_corner / voxel) * voxel and every occupied voxel is replaced by the barycentre of its points.
//
// The reference bins through a std::unordered_map and emits the voxels in that container's iteration order, which is a
// property of libstdc++'s hash table, not of the algorithm.  Here the voxel grid of a batch element is DENSE in device
// memory (3 m / 0.05 m = 60^3 cells for 3DMatch, 267 x 267 x 14 for KITTI: megabytes), points are accumulated with
// atomics, and the occupied cells are compacted in ascending cell index: a deterministic order.  Parity is therefore
// defined on the SET of barycentres (tests sort both sides).
// Arithmetic: the cell of a point reproduces the reference's fp32 sequence bit for bit
// (origin = floor(min * float(1/voxel)) * voxel; i = floor((p - origin) / voxel), cloud.h:88-108,
// grid_subsampling_cpu.cpp:9-36), so the number of voxels and their populations are exact.  Sums are accumulated in
// fp64 (order-independent to 1e-16, so the result does not depend on the atomics' order) and rounded once; the
// reference adds in fp32 in input order: barycentres agree to a few fp32 ulp (tests: 2e-6 relative to the extent).
#include "geotx_common.cuh"

namespace geotx {

struct GsCell {
    double sx, sy, sz;
    int count;
    int rank;  // position among the occupied cells of this batch element (filled by the compaction)
};

struct GsGrid {            // per batch element, on the device
    float origin[3];
    unsigned long long nx, ny, nz, cells;
    int overflow;          // cells > capacity
};

struct GsWorkspace {
    GsGrid* grid;                    // [1]
    float* minmax;                   // [6]
    unsigned int* block_counts;      // [capacity / GS_SCAN_BLOCK + 1]
    unsigned long long* out_base;    // [1] barycentres written by the previous batch elements
    GsCell* cells;                   // [capacity]
};

constexpr int GS_SCAN_BLOCK = 1024;

__host__ inline size_t gs_align(size_t x) { return (x + 255) & ~(size_t)255; }
__host__ inline size_t gs_layout(int64_t capacity, char* base, GsWorkspace* w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += gs_align(bytes); return p; };
    char* a = take(sizeof(GsGrid));
    char* b = take(6 * sizeof(float));
    char* c = take(((size_t)capacity / GS_SCAN_BLOCK + 2) * sizeof(unsigned int));
    char* d = take(sizeof(unsigned long long));
    char* e = take((size_t)capacity * sizeof(GsCell));
    if (w) {
        w->grid = (GsGrid*)a;
        w->minmax = (float*)b;
        w->block_counts = (unsigned int*)c;
        w->out_base = (unsigned long long*)d;
        w->cells = (GsCell*)e;
    }
    return off;
}

// cloud.cpp:4-40 min_point / max_point, then grid_subsampling_cpu.cpp:9-20 (one CTA)
__global__ void __launch_bounds__(1024)
gs_bounds_kernel(const float* __restrict__ points, int n, float voxel, long long capacity, GsWorkspace w) {
    __shared__ float red[6][32];
    float lo[3] = {__int_as_float(0x7f800000), __int_as_float(0x7f800000), __int_as_float(0x7f800000)};
    float hi[3] = {-__int_as_float(0x7f800000), -__int_as_float(0x7f800000), -__int_as_float(0x7f800000)};
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const float v = points[3 * (size_t)i + d];
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
        if ((threadIdx.x & 31) == 0) { red[d][threadIdx.x >> 5] = lo[d]; red[3 + d][threadIdx.x >> 5] = hi[d]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const int nw = blockDim.x >> 5;
        const float inv = (float)(1.0 / (double)voxel);  // "1. / voxel_size" is a double, narrowed by operator*(PointXYZ, float)
        unsigned long long dims[3];
        for (int d = 0; d < 3; ++d) {
            float l = red[d][0], h = red[3 + d][0];
            for (int k = 1; k < nw; ++k) { l = fminf(l, red[d][k]); h = fmaxf(h, red[3 + d][k]); }
            const float origin = __fmul_rn(floorf(__fmul_rn(l, inv)), voxel);              // :11
            w.grid->origin[d] = origin;
            dims[d] = (unsigned long long)(floorf(__fdiv_rn(__fsub_rn(h, origin), voxel)) + 1.0f);  // :13-20
        }
        // z extent: the reference never needs it (its map is sparse); the dense grid does.  A point on the upper face
        // of the bounding box can land one cell further than floor((max - origin) / voxel) only through rounding of
        // the same expression, which is what dims[] already evaluates: no slack needed.
        w.grid->nx = dims[0];
        w.grid->ny = dims[1];
        w.grid->nz = dims[2];
        const unsigned long long cells = dims[0] * dims[1] * dims[2];
        w.grid->cells = cells;
        w.grid->overflow = (dims[0] > 0xFFFFFu || dims[1] > 0xFFFFFu || dims[2] > 0xFFFFFu || cells > (unsigned long long)capacity) ? 1 : 0;
    }
}

__global__ void __launch_bounds__(256) gs_clear_kernel(GsWorkspace w) {
    if (w.grid->overflow) return;
    const unsigned long long cells = w.grid->cells;
    for (unsigned long long c = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; c < cells;
         c += (unsigned long long)gridDim.x * blockDim.x) {
        GsCell z;
        z.sx = 0.0; z.sy = 0.0; z.sz = 0.0; z.count = 0; z.rank = 0;
        w.cells[c] = z;
    }
}

// grid_subsampling_cpu.cpp:28-43
__global__ void __launch_bounds__(256)
gs_accumulate_kernel(const float* __restrict__ points, int n, float voxel, GsWorkspace w) {
    if (w.grid->overflow) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = points[3 * (size_t)i], y = points[3 * (size_t)i + 1], z = points[3 * (size_t)i + 2];
    const unsigned long long ix = (unsigned long long)floorf(__fdiv_rn(__fsub_rn(x, w.grid->origin[0]), voxel));
    const unsigned long long iy = (unsigned long long)floorf(__fdiv_rn(__fsub_rn(y, w.grid->origin[1]), voxel));
    const unsigned long long iz = (unsigned long long)floorf(__fdiv_rn(__fsub_rn(z, w.grid->origin[2]), voxel));
    GsCell* c = w.cells + (ix + w.grid->nx * iy + w.grid->nx * w.grid->ny * iz);  // :36
    atomicAdd(&c->sx, (double)x);
    atomicAdd(&c->sy, (double)y);
    atomicAdd(&c->sz, (double)z);
    atomicAdd(&c->count, 1);
}

// occupied cells per tile of GS_SCAN_BLOCK cells (persistent CTAs walk the tiles the grid really has)
__global__ void __launch_bounds__(GS_SCAN_BLOCK) gs_count_kernel(GsWorkspace w) {
    if (w.grid->overflow) return;
    const unsigned long long cells = w.grid->cells;
    for (unsigned long long tile = blockIdx.x; tile * GS_SCAN_BLOCK < cells; tile += gridDim.x) {
        const unsigned long long c = tile * GS_SCAN_BLOCK + threadIdx.x;
        const int n = __syncthreads_count(c < cells && w.cells[c].count > 0);
        if (threadIdx.x == 0) w.block_counts[tile] = (unsigned int)n;
    }
}

// exclusive scan of the block counts (one CTA); s_lengths[b] = total
__global__ void __launch_bounds__(1024) gs_scan_kernel(int nblocks, GsWorkspace w, long long* __restrict__ s_length) {
    __shared__ unsigned int warp_sum[32];
    __shared__ unsigned int carry_s;
    if (w.grid->overflow) {
        if (threadIdx.x == 0) *s_length = -1;
        return;
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int used = (int)((w.grid->cells + GS_SCAN_BLOCK - 1) / GS_SCAN_BLOCK);
    for (int base = 0; base < min(nblocks, used); base += 1024) {
        const int i = base + threadIdx.x;
        const unsigned int c = i < used ? w.block_counts[i] : 0u;
        unsigned int incl = c;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const unsigned int o = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += o;
        }
        if (lane == 31) warp_sum[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            unsigned int s = warp_sum[lane];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const unsigned int o = __shfl_up_sync(0xffffffffu, s, off);
                if (lane >= off) s += o;
            }
            warp_sum[lane] = s;
        }
        __syncthreads();
        const unsigned int before = carry_s + (wid ? warp_sum[wid - 1] : 0u) + incl - c;
        if (i < used) w.block_counts[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = before + c;
        __syncthreads();
    }
    if (threadIdx.x == 0) *s_length = (long long)carry_s;
}

// grid_subsampling_cpu.cpp:45-48: barycentre = sum * float(1 / count), written in ascending cell order
__global__ void __launch_bounds__(GS_SCAN_BLOCK) gs_emit_kernel(GsWorkspace w, float* __restrict__ s_points) {
    __shared__ unsigned int warp_sum[32];
    if (w.grid->overflow) return;
    const unsigned long long cells = w.grid->cells;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (unsigned long long tile = blockIdx.x; tile * GS_SCAN_BLOCK < cells; tile += gridDim.x) {
        const unsigned long long c = tile * GS_SCAN_BLOCK + threadIdx.x;
        const bool occ = c < cells && w.cells[c].count > 0;
        const unsigned int ballot = __ballot_sync(0xffffffffu, occ);
        __syncthreads();  // warp_sum of the previous tile has been read
        if (lane == 0) warp_sum[wid] = __popc(ballot);
        __syncthreads();
        if (wid == 0) {
            unsigned int s = warp_sum[lane];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const unsigned int o = __shfl_up_sync(0xffffffffu, s, off);
                if (lane >= off) s += o;
            }
            warp_sum[lane] = s;  // inclusive
        }
        __syncthreads();
        if (!occ) continue;
        const unsigned long long rank = *w.out_base + w.block_counts[tile] + (wid ? warp_sum[wid - 1] : 0u) +
                                        __popc(ballot & ((1u << lane) - 1u));
        const GsCell cell = w.cells[c];
        const float inv = (float)(1.0 / (double)cell.count);
        s_points[3 * rank] = __fmul_rn((float)cell.sx, inv);
        s_points[3 * rank + 1] = __fmul_rn((float)cell.sy, inv);
        s_points[3 * rank + 2] = __fmul_rn((float)cell.sz, inv);
    }
}

__global__ void gs_advance_kernel(GsWorkspace w, const long long* __restrict__ s_length) {
    if (*s_length > 0) *w.out_base += (unsigned long long)*s_length;
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
#include "host_util.h"
using namespace geotx;

extern "C" int64_t geotx_grid_subsample_workspace_bytes(int64_t max_cells) {
    if (max_cells <= 0 || max_cells > ((int64_t)1 << 31)) return 0;
    return (int64_t)gs_layout(max_cells, nullptr, nullptr);
}

extern "C" int geotx_grid_subsample(const float* points, const int64_t* host_lengths, int batch, float voxel_size,
                                    float* s_points, int64_t* s_lengths, void* workspace, int64_t max_cells, void* stream) {
    GEOTX_REQUIRE(points && host_lengths && s_points && s_lengths && workspace, "geotx_grid_subsample: null pointer");
    GEOTX_REQUIRE(batch > 0 && voxel_size > 0.f, "geotx_grid_subsample: bad batch=%d or voxel_size=%g", batch, voxel_size);
    GEOTX_REQUIRE(max_cells > 0 && max_cells <= ((int64_t)1 << 31), "geotx_grid_subsample: max_cells=%lld outside (0, 2^31]",
                  (long long)max_cells);
    GEOTX_REQUIRE(((uintptr_t)workspace & 255) == 0, "geotx_grid_subsample: workspace must be 256-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    GsWorkspace w;
    gs_layout(max_cells, (char*)workspace, &w);
    int rc = check_cuda(cudaMemsetAsync(w.out_base, 0, sizeof(unsigned long long), st), "geotx_grid_subsample: memset");
    if (rc) return rc;
    const int nblocks = (int)((max_cells + GS_SCAN_BLOCK - 1) / GS_SCAN_BLOCK);
    int64_t start = 0;
    const int sms = sm_count();
    for (int b = 0; b < batch; ++b) {
        const int64_t n = host_lengths[b];
        GEOTX_REQUIRE(n > 0 && n < 2147483647LL, "geotx_grid_subsample: batch element %d has %lld points (the reference "
                      "reads points[0] of every element, cloud.cpp:5)", b, (long long)n);
        const float* p = points + 3 * start;
        gs_bounds_kernel<<<1, 1024, 0, st>>>(p, (int)n, voxel_size, (long long)max_cells, w);
        gs_clear_kernel<<<sms * 8, 256, 0, st>>>(w);
        gs_accumulate_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p, (int)n, voxel_size, w);
        gs_count_kernel<<<sms * 2, GS_SCAN_BLOCK, 0, st>>>(w);
        gs_scan_kernel<<<1, 1024, 0, st>>>(nblocks, w, (long long*)s_lengths + b);
        gs_emit_kernel<<<sms * 2, GS_SCAN_BLOCK, 0, st>>>(w, s_points);
        gs_advance_kernel<<<1, 1, 0, st>>>(w, (const long long*)s_lengths + b);
        start += n;
    }
    return launch_status("geotx_grid_subsample");
}
