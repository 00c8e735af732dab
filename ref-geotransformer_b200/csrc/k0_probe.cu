// Diagnostic (no reference counterpart): the legacy tensor path's peak on this GPU, measured the way the kernels that use
// it run -- mma.sync.m16n8k16 (f16 operands, fp32 accumulation), 16 independent accumulators per warp, operands in
// registers.  bench.py times it with CUDA events and reports K3r's fraction of THIS pipe beside its fraction of the
// measured tcgen05 (cuBLAS bf16) rate.
#include "geotx_common.cuh"
#include "host_util.h"

namespace geotx {
__global__ void __launch_bounds__(256) mma_sync_probe_kernel(int iters, float* __restrict__ out) {
    unsigned a[4], b[4][2];
    float d[16][4] = {};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        a[i] = 0x3c003c00u + threadIdx.x + i;
        b[i][0] = 0x38003800u + i;
        b[i][1] = 0x34003400u + i;
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int t = 0; t < 16; ++t)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(d[t][0]), "+f"(d[t][1]), "+f"(d[t][2]), "+f"(d[t][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[t & 3][0]), "r"(b[t & 3][1]));
    }
    float s = 0.f;
#pragma unroll
    for (int t = 0; t < 16; ++t) s += d[t][0] + d[t][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace geotx

using namespace geotx;

// ctas x 8 warps x iters x 16 MMAs of 4096 flop each; out: f32 [ctas * 256]
extern "C" int geotx_mma_sync_probe(int ctas, int iters, float* out, void* stream) {
    GEOTX_REQUIRE(ctas > 0 && iters > 0 && out, "geotx_mma_sync_probe: bad arguments");
    mma_sync_probe_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>(iters, out);
    return launch_status("geotx_mma_sync_probe");
}
