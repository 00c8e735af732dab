// C entry points over the REFERENCE's own CPU collate operators, compiled from the reference sources where they lie
// (never copied into this repo):
//   /root/reference/geotransformer/extensions/cpu/radius_neighbors/radius_neighbors_cpu.cpp
//   /root/reference/geotransformer/extensions/cpu/grid_subsampling/grid_subsampling_cpu.cpp
//   /root/reference/geotransformer/extensions/extra/cloud/cloud.cpp   (+ the vendored nanoflann.hpp)
// These three files are torch-free (the tensor marshalling lives in radius_neighbors.cpp / grid_subsampling.cpp, which
// this shim replaces with plain arrays, mirroring cpu/radius_neighbors/radius_neighbors.cpp:3-76 and
// cpu/grid_subsampling/grid_subsampling.cpp).
//
// TEST INFRASTRUCTURE ONLY: built by oracle/build_ref.sh into oracle/_ref/libgeotx_ref_cpu.so (git-ignored), used by
// tests/ as the checker and by bench.py's CPU legs as the timed reference.  The product never loads it.
#include <cstdint>
#include <cstring>
#include <vector>

#include "geotransformer/extensions/cpu/grid_subsampling/grid_subsampling_cpu.h"
#include "geotransformer/extensions/cpu/radius_neighbors/radius_neighbors_cpu.h"

static std::vector<long> g_neighbors;   // results are kept until the next call (single-threaded test use)
static std::vector<PointXYZ> g_sub_points;
static std::vector<long> g_sub_lengths;

static std::vector<PointXYZ> to_points(const float* p, int64_t n) {
    std::vector<PointXYZ> v((size_t)n);
    std::memcpy(reinterpret_cast<void*>(v.data()), p, sizeof(float) * 3 * (size_t)n);  // radius_neighbors.cpp:37-48
    return v;
}

extern "C" {

// radius_neighbors.cpp:3-76.  Returns max_count (the width of the result); geotx_ref_radius_neighbors_copy fetches it.
int64_t geotx_ref_radius_neighbors(const float* q_points, int64_t nq, const float* s_points, int64_t ns,
                                   const int64_t* q_lengths, const int64_t* s_lengths, int batch, float radius) {
    std::vector<PointXYZ> q = to_points(q_points, nq), s = to_points(s_points, ns);
    std::vector<long> ql(q_lengths, q_lengths + batch), sl(s_lengths, s_lengths + batch);
    g_neighbors.clear();
    radius_neighbors_cpu(q, s, ql, sl, g_neighbors, radius);
    return nq > 0 ? (int64_t)(g_neighbors.size() / (size_t)nq) : 0;
}

void geotx_ref_radius_neighbors_copy(int64_t* out) {
    for (size_t i = 0; i < g_neighbors.size(); ++i) out[i] = (int64_t)g_neighbors[i];
}

// grid_subsampling.cpp.  Returns the number of subsampled points.
int64_t geotx_ref_grid_subsampling(const float* points, int64_t n, const int64_t* lengths, int batch, float voxel_size) {
    std::vector<PointXYZ> p = to_points(points, n);
    std::vector<long> l(lengths, lengths + batch);
    g_sub_points.clear();
    g_sub_lengths.clear();
    grid_subsampling_cpu(p, g_sub_points, l, g_sub_lengths, voxel_size);
    return (int64_t)g_sub_points.size();
}

void geotx_ref_grid_subsampling_copy(float* s_points, int64_t* s_lengths) {
    std::memcpy(s_points, reinterpret_cast<const void*>(g_sub_points.data()), sizeof(float) * 3 * g_sub_points.size());
    for (size_t i = 0; i < g_sub_lengths.size(); ++i) s_lengths[i] = (int64_t)g_sub_lengths[i];
}

}  // extern "C"
