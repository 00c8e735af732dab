// K6: point-to-node partition -- every fine point goes to its nearest node (superpoint), every
// node keeps its `point_limit` nearest OWN points in ascending distance.
//
// Replaces /root/reference/geotransformer/modules/ops/pointcloud_partition.py:61-107 (called twice per pair from
// experiments/*/model.py:98-103) and the patch gather behind it (model.py:105-108).  The reference materialises the
// (M,N) fp32 distance matrix, an (M,N) bool matrix, a masked copy filled with 1e12 and a full top-k over N per node.
// Here nothing of size M x N exists:
//   1. assign   one pass over (point, node) pairs in registers, nodes staged in shared memory; arg-min per point;
//               the winning (distance, point) pair becomes a 64-bit sort key; per-node histogram by atomics
//   2. offsets  exclusive scan of the histogram (one CTA); node_sizes, node_masks
//   3. scatter  counting sort of the keys into per-node buckets (order inside a bucket is arbitrary ...)
//   4. select   ... and is fixed here: one CTA per node bitonic-sorts its bucket in shared memory (chunked merge that
//               keeps the best point_limit keys when a bucket is larger than the buffer) and writes indices, masks and
//               the gathered patch points.  Keys are unique (they contain the point index), so the result does not
//               depend on the scatter order: the output is deterministic.
// Distances reproduce the reference's fp32 op order bit for bit (geotx_common.cuh: sqdist_ref with x = node,
// y = point, exactly pairwise_distance(nodes, points)); equal distances are ordered by the smaller index
// (torch.min / torch.topk leave that order unspecified).
#include "geotx_common.cuh"
#include <mutex>

namespace geotx {

constexpr int P2N_POINTS_PER_BLOCK = 64;
constexpr int P2N_SLICES = 4;            // node slices per point: 64 x 4 = 256 threads
constexpr int P2N_NODE_TILE = 2048;      // nodes staged per shared-memory tile (32 KB)

struct P2NWorkspace {
    unsigned long long* keys;     // [N] (d2 bits << 32) | point
    unsigned long long* bucket;   // [N] keys grouped by node
    unsigned int* counts;         // [M]
    unsigned int* cursor;         // [M]
    unsigned int* offsets;        // [M+1]
    int* owner;                   // [N] node of each point (32-bit copy for the scatter)
};

__host__ __device__ inline size_t p2n_align(size_t x) { return (x + 255) & ~(size_t)255; }

__host__ inline size_t p2n_layout(int N, int M, char* base, P2NWorkspace* w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += p2n_align(bytes); return p; };
    char* a = take((size_t)N * 8);
    char* b = take((size_t)N * 8);
    char* c = take((size_t)M * 4 * 2);  // counts and cursor are cleared by one memset
    char* d = take((size_t)(M + 1) * 4);
    char* e = take((size_t)N * 4);
    if (w) {
        w->keys = (unsigned long long*)a;
        w->bucket = (unsigned long long*)b;
        w->counts = (unsigned int*)c;
        w->cursor = w->counts + M;
        w->offsets = (unsigned int*)d;
        w->owner = (int*)e;
    }
    return off;
}

// pointcloud_partition.py:82-84: sq_dist_mat = pairwise_distance(nodes, points); point_to_node = argmin over nodes.
__global__ void __launch_bounds__(P2N_POINTS_PER_BLOCK * P2N_SLICES)
p2n_assign_kernel(const float* __restrict__ points, int N, const float* __restrict__ nodes, int M,
                  long long* __restrict__ point_to_node, P2NWorkspace w) {
    __shared__ float4 tile[P2N_NODE_TILE];
    __shared__ float red_d[P2N_SLICES][P2N_POINTS_PER_BLOCK];
    __shared__ int red_i[P2N_SLICES][P2N_POINTS_PER_BLOCK];
    const int pl = threadIdx.x % P2N_POINTS_PER_BLOCK;   // a warp = 32 consecutive points of one slice
    const int slice = threadIdx.x / P2N_POINTS_PER_BLOCK;
    const int p = blockIdx.x * P2N_POINTS_PER_BLOCK + pl;
    float4 pt = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p < N) {
        const float x = points[3 * (size_t)p], y = points[3 * (size_t)p + 1], z = points[3 * (size_t)p + 2];
        pt = make_float4(x, y, z, sq_norm_ref(x, y, z));
    }
    float best = __int_as_float(0x7f800000);
    int best_i = 0x7fffffff;
    for (int t0 = 0; t0 < M; t0 += P2N_NODE_TILE) {
        const int cnt = min(P2N_NODE_TILE, M - t0);
        __syncthreads();
        for (int q = threadIdx.x; q < cnt; q += blockDim.x) {
            const float x = nodes[3 * (size_t)(t0 + q)], y = nodes[3 * (size_t)(t0 + q) + 1],
                        z = nodes[3 * (size_t)(t0 + q) + 2];
            tile[q] = make_float4(x, y, z, sq_norm_ref(x, y, z));
        }
        __syncthreads();
        // all lanes of a warp read the same node: shared-memory broadcast
#pragma unroll 4
        for (int q = slice; q < cnt; q += P2N_SLICES) {
            const float d = sqdist_ref(tile[q], pt);  // x = node, y = point
            if (d < best) { best = d; best_i = t0 + q; }  // q only grows within a thread: first minimum kept
        }
    }
    red_d[slice][pl] = best;
    red_i[slice][pl] = best_i;
    __syncthreads();
    if (slice == 0 && p < N) {
#pragma unroll
        for (int s = 1; s < P2N_SLICES; ++s) {
            const float d = red_d[s][pl];
            const int i = red_i[s][pl];
            if (d < best || (d == best && i < best_i)) { best = d; best_i = i; }
        }
        point_to_node[p] = (long long)best_i;
        w.owner[p] = best_i;
        w.keys[p] = ((unsigned long long)__float_as_uint(best) << 32) | (unsigned int)p;  // d2 >= 0: bits are ordered
        atomicAdd(w.counts + best_i, 1u);
    }
}

// Exclusive scan of the per-node counts by one CTA; node_sizes / node_masks (pointcloud_partition.py:85-86, :101-104).
__global__ void __launch_bounds__(1024)
p2n_offsets_kernel(int M, P2NWorkspace w, long long* __restrict__ node_sizes, uint8_t* __restrict__ node_masks) {
    __shared__ unsigned int warp_sum[32];
    __shared__ unsigned int carry_s;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < M; base += 1024) {
        const int i = base + threadIdx.x;
        const unsigned int c = i < M ? w.counts[i] : 0u;
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
            warp_sum[lane] = s;  // inclusive over warps
        }
        __syncthreads();
        const unsigned int carry = carry_s;
        const unsigned int before = carry + (wid ? warp_sum[wid - 1] : 0u) + incl - c;
        if (i < M) {
            w.offsets[i] = before;
            if (node_sizes) node_sizes[i] = (long long)c;
            node_masks[i] = c ? 1 : 0;
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = before + c;
        __syncthreads();
    }
    if (threadIdx.x == 0) w.offsets[M] = carry_s;
}

__global__ void __launch_bounds__(256) p2n_scatter_kernel(int N, P2NWorkspace w) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= N) return;
    const int node = w.owner[p];
    const unsigned int pos = w.offsets[node] + atomicAdd(w.cursor + node, 1u);
    w.bucket[pos] = w.keys[p];
}

// Ascending bitonic sort of buf[0, len) (len a power of two) by the whole CTA.
__device__ __forceinline__ void bitonic_sort(unsigned long long* buf, int len) {
    for (int k = 2; k <= len; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < (len >> 1); t += blockDim.x) {
                const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // insert a 0 bit at position log2(j)
                const int hi = lo | j;
                const bool up = (lo & k) == 0;
                const unsigned long long a = buf[lo], b = buf[hi];
                if ((a > b) == up) { buf[lo] = b; buf[hi] = a; }
            }
            __syncthreads();
        }
    }
}

// pointcloud_partition.py:88-98: per node the point_limit nearest points among those assigned to it, ascending;
// slots beyond the node's size hold the sentinel index N and mask 0.  model.py:105-108: the gathered patch points
// (the sentinel row of the padded cloud is all zeros).
__global__ void __launch_bounds__(128)
p2n_select_kernel(const float* __restrict__ points, int N, int K, int sort_n, P2NWorkspace w,
                  long long* __restrict__ knn_indices, uint8_t* __restrict__ knn_masks, float* __restrict__ knn_points) {
    extern __shared__ unsigned long long buf[];  // sort_n keys
    const int node = blockIdx.x;
    const unsigned int begin = w.offsets[node];
    const int count = (int)(w.offsets[node + 1] - begin);
    int filled = 0;  // sorted keys kept from the previous chunks, at buf[0, filled)
    int base = 0;
    while (base < count) {
        const int take = min(count - base, sort_n - filled);
        for (int t = threadIdx.x; t < take; t += blockDim.x) buf[filled + t] = w.bucket[begin + base + t];
        const int len = filled + take;
        int p2 = 1;
        while (p2 < len) p2 <<= 1;
        for (int t = len + threadIdx.x; t < p2; t += blockDim.x) buf[t] = ~0ull;
        __syncthreads();
        bitonic_sort(buf, p2);
        filled = min(K, len);
        base += take;
    }
    __syncthreads();
    const int valid = min(count, K);
    for (int r = threadIdx.x; r < K; r += blockDim.x) {
        const size_t o = (size_t)node * K + r;
        if (r < valid) {
            const unsigned int p = (unsigned int)(buf[r] & 0xffffffffull);
            knn_indices[o] = (long long)p;
            knn_masks[o] = 1;
            if (knn_points) {
                knn_points[3 * o] = points[3 * (size_t)p];
                knn_points[3 * o + 1] = points[3 * (size_t)p + 1];
                knn_points[3 * o + 2] = points[3 * (size_t)p + 2];
            }
        } else {
            knn_indices[o] = (long long)N;  // :98 masked_fill_(~node_knn_masks, points.shape[0])
            knn_masks[o] = 0;
            if (knn_points) { knn_points[3 * o] = 0.f; knn_points[3 * o + 1] = 0.f; knn_points[3 * o + 2] = 0.f; }
        }
    }
}

}  // namespace geotx

// ---- C ABI --------------------------------------------------------------------------------------------------
#include "host_util.h"
using namespace geotx;

static int p2n_sort_n(int point_limit) {
    int s = 256;
    while (s < 2 * point_limit) s <<= 1;
    return s;
}

extern "C" int64_t geotx_p2n_workspace_bytes(int num_points, int num_nodes, int point_limit) {
    if (num_points <= 0 || num_nodes <= 0 || point_limit <= 0 || point_limit > GEOTX_MAX_POINT_LIMIT) return 0;
    return (int64_t)p2n_layout(num_points, num_nodes, nullptr, nullptr);
}

extern "C" int geotx_point_to_node_partition(const float* points, int num_points, const float* nodes, int num_nodes,
                                             int point_limit, int64_t* point_to_node, int64_t* node_sizes,
                                             uint8_t* node_masks, int64_t* node_knn_indices, uint8_t* node_knn_masks,
                                             float* node_knn_points, void* workspace, void* stream) {
    GEOTX_REQUIRE(points && nodes && point_to_node && node_masks && node_knn_indices && node_knn_masks && workspace,
                  "geotx_point_to_node_partition: null pointer");
    GEOTX_REQUIRE(num_points > 0 && num_nodes > 0, "geotx_point_to_node_partition: empty input (points=%d nodes=%d)",
                  num_points, num_nodes);
    // torch.topk raises for k > N (pointcloud_partition.py:93); so do we
    GEOTX_REQUIRE(point_limit >= 1 && point_limit <= num_points,
                  "geotx_point_to_node_partition: selected index k out of range (point_limit=%d, points=%d)", point_limit,
                  num_points);
    GEOTX_REQUIRE(point_limit <= GEOTX_MAX_POINT_LIMIT, "geotx_point_to_node_partition: point_limit=%d > %d", point_limit,
                  GEOTX_MAX_POINT_LIMIT);
    GEOTX_REQUIRE(((uintptr_t)workspace & 255) == 0, "geotx_point_to_node_partition: workspace must be 256-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    P2NWorkspace w;
    p2n_layout(num_points, num_nodes, (char*)workspace, &w);
    int rc = check_cuda(cudaMemsetAsync(w.counts, 0, (size_t)num_nodes * 8, st), "geotx_point_to_node_partition: memset");
    if (rc) return rc;
    p2n_assign_kernel<<<(num_points + P2N_POINTS_PER_BLOCK - 1) / P2N_POINTS_PER_BLOCK,
                        P2N_POINTS_PER_BLOCK * P2N_SLICES, 0, st>>>(points, num_points, nodes, num_nodes,
                                                                    (long long*)point_to_node, w);
    p2n_offsets_kernel<<<1, 1024, 0, st>>>(num_nodes, w, (long long*)node_sizes, node_masks);
    p2n_scatter_kernel<<<(num_points + 255) / 256, 256, 0, st>>>(num_points, w);
    const int sort_n = p2n_sort_n(point_limit);
    const size_t smem = (size_t)sort_n * 8;
    if (smem > 48 * 1024) {
        rc = check_cuda(cudaFuncSetAttribute(p2n_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                        "geotx_point_to_node_partition: shared memory");
        if (rc) return rc;
    }
    p2n_select_kernel<<<num_nodes, 128, smem, st>>>(points, num_points, point_limit, sort_n, w,
                                                    (long long*)node_knn_indices, node_knn_masks, node_knn_points);
    return launch_status("geotx_point_to_node_partition");
}

// ---- allocation-free variant (no counterpart in the reference; see include/geotx_b200.h) --------------------------
// The workspace of the entry point above lives in a file-scope device array instead of caller memory.  Every GPU that
// loads the library holds these GEOTX_P2N_STATIC_SCRATCH_BYTES until the process exits.  In host code the array's name
// is only a handle: its device address is looked up once per device with cudaGetSymbolAddress (that device current) and
// cached; the kernels receive that address.  All calls on one device share the array, so calls that overlap on the GPU
// (different streams) race; calls on one stream are ordered and each call clears the counters it needs.
static __device__ __align__(256) unsigned char g_p2n_static_scratch[GEOTX_P2N_STATIC_SCRATCH_BYTES];

constexpr int P2N_MAX_DEVICES = 64;
static void* g_p2n_static_addr[P2N_MAX_DEVICES];  // device address of g_p2n_static_scratch, per device (0 = not yet)
static std::mutex g_p2n_static_mutex;

static int p2n_static_scratch_address(void** out) {
    int dev = -1;
    int rc = check_cuda(cudaGetDevice(&dev), "geotx_point_to_node_partition_static: cudaGetDevice");
    if (rc) return rc;
    GEOTX_REQUIRE(dev >= 0 && dev < P2N_MAX_DEVICES, "geotx_point_to_node_partition_static: device %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_p2n_static_mutex);
    if (!g_p2n_static_addr[dev]) {
        void* ptr = nullptr;
        rc = check_cuda(cudaGetSymbolAddress(&ptr, g_p2n_static_scratch),
                        "geotx_point_to_node_partition_static: cudaGetSymbolAddress");
        if (rc) return rc;
        GEOTX_REQUIRE(ptr && ((uintptr_t)ptr & 255) == 0,
                      "geotx_point_to_node_partition_static: scratch symbol resolved to a bad address");
        g_p2n_static_addr[dev] = ptr;
    }
    *out = g_p2n_static_addr[dev];
    return GEOTX_OK;
}

extern "C" int64_t geotx_p2n_static_scratch_bytes(void) { return (int64_t)GEOTX_P2N_STATIC_SCRATCH_BYTES; }

extern "C" int geotx_point_to_node_partition_static(const float* points, int num_points, const float* nodes,
                                                    int num_nodes, int point_limit, int64_t* point_to_node,
                                                    int64_t* node_sizes, uint8_t* node_masks, int64_t* node_knn_indices,
                                                    uint8_t* node_knn_masks, float* node_knn_points, void* stream) {
    GEOTX_REQUIRE(points && nodes && point_to_node && node_masks && node_knn_indices && node_knn_masks,
                  "geotx_point_to_node_partition_static: null pointer");
    GEOTX_REQUIRE(num_points > 0 && num_nodes > 0,
                  "geotx_point_to_node_partition_static: empty input (points=%d nodes=%d)", num_points, num_nodes);
    GEOTX_REQUIRE(point_limit >= 1 && point_limit <= num_points,
                  "geotx_point_to_node_partition_static: selected index k out of range (point_limit=%d, points=%d)",
                  point_limit, num_points);
    GEOTX_REQUIRE(point_limit <= GEOTX_MAX_POINT_LIMIT, "geotx_point_to_node_partition_static: point_limit=%d > %d",
                  point_limit, GEOTX_MAX_POINT_LIMIT);
    const size_t need = p2n_layout(num_points, num_nodes, nullptr, nullptr);
    GEOTX_REQUIRE(need <= (size_t)GEOTX_P2N_STATIC_SCRATCH_BYTES,
                  "geotx_point_to_node_partition_static: points=%d nodes=%d need %zu bytes of scratch, the built-in "
                  "array holds %d; use geotx_point_to_node_partition with a caller workspace",
                  num_points, num_nodes, need, (int)GEOTX_P2N_STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    int rc = p2n_static_scratch_address(&scratch);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    P2NWorkspace w;
    p2n_layout(num_points, num_nodes, (char*)scratch, &w);
    // the array keeps the previous call's contents: the histogram and the scatter cursors must start at zero
    rc = check_cuda(cudaMemsetAsync(w.counts, 0, (size_t)num_nodes * 8, st), "geotx_point_to_node_partition_static: memset");
    if (rc) return rc;
    p2n_assign_kernel<<<(num_points + P2N_POINTS_PER_BLOCK - 1) / P2N_POINTS_PER_BLOCK,
                        P2N_POINTS_PER_BLOCK * P2N_SLICES, 0, st>>>(points, num_points, nodes, num_nodes,
                                                                    (long long*)point_to_node, w);
    p2n_offsets_kernel<<<1, 1024, 0, st>>>(num_nodes, w, (long long*)node_sizes, node_masks);
    p2n_scatter_kernel<<<(num_points + 255) / 256, 256, 0, st>>>(num_points, w);
    const int sort_n = p2n_sort_n(point_limit);
    const size_t smem = (size_t)sort_n * 8;
    if (smem > 48 * 1024) {
        rc = check_cuda(cudaFuncSetAttribute(p2n_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                        "geotx_point_to_node_partition_static: shared memory");
        if (rc) return rc;
    }
    p2n_select_kernel<<<num_nodes, 128, smem, st>>>(points, num_points, point_limit, sort_n, w,
                                                    (long long*)node_knn_indices, node_knn_masks, node_knn_points);
    return launch_status("geotx_point_to_node_partition_static");
}
