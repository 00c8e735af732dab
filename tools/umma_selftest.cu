// Stand-alone self-test of the tcgen05 plumbing used by k2_gse_embed.cu (descriptors, 128B swizzle, TMEM layout,
// CTA-pair MMA with remote mbarrier arrivals and multicast commit).  Not part of the product library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I ref-geotransformer_b200/csrc tools/umma_selftest.cu -o /tmp/umma_selftest
//   umma_selftest one | pair
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "sm100_ptx.cuh"

using namespace geotx;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int KC = 64;  // one 128-byte swizzled row

// A: [rows][64] bf16 row-major in global; writes the swizzled K-major tile (rows x 128 B) at smem `dst`
__device__ void fill_swizzled(uint8_t* dst, const __nv_bfloat16* src, int rows, int tid, int nthreads) {
    for (int idx = tid; idx < rows * 8; idx += nthreads) {
        int r = idx >> 3, g = idx & 7;
        uint4 v = *(const uint4*)(src + (size_t)r * KC + g * 8);
        *(uint4*)(dst + r * 128 + ((g ^ (r & 7)) << 4)) = v;
    }
}

// ---- variant 1: one CTA, M=128, N=256, K=64 ----
__global__ void __launch_bounds__(128) one_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int N) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                // 128 x 128 B
    uint8_t* sB = smem + 16384;        // N x 128 B
    uint64_t* bar = (uint64_t*)(sB + 256 * 128);
    uint32_t* s_tmem = (uint32_t*)(bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { ptx::mbar_init(smem_u32(bar), 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc(smem_u32(s_tmem), 256);
    fill_swizzled(sA, A, 128, threadIdx.x, 128);
    fill_swizzled(sB, B, N, threadIdx.x, 128);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = *s_tmem;
    if (threadIdx.x == 0) {
        const uint32_t idesc = ptx::umma_idesc_bf16(128, N);
        const uint64_t ad = ptx::umma_desc_sw128(smem_u32(sA)), bd = ptx::umma_desc_sw128(smem_u32(sB));
        for (int ks = 0; ks < 4; ++ks) ptx::umma_bf16(tmem, ad + 2 * ks, bd + 2 * ks, idesc, ks > 0);
        ptx::umma_commit(smem_u32(bar));
    }
    ptx::mbar_wait(smem_u32(bar), 0);
    ptx::tc_fence_after_sync();
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v[16];
        ptx::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
        ptx::tmem_ld_wait();
        for (int c = 0; c < 16; ++c) D[(size_t)row * N + c0 + c] = __uint_as_float(v[c]);
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tmem, 256);
}

// ---- variant 2: CTA pair, M=256 (128 rows per CTA), N=256 (128 B-rows per CTA), K=64 ----
// A: [256][64]; B: [256][64]; CTA r fills A rows 128r.., B rows 128r..; D: [256][256]
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128)
pair_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int N) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;
    uint8_t* sB = smem + 16384;
    uint64_t* bars = (uint64_t*)(sB + 128 * 128);  // [0] full (leader: 8 warp arrivals), [1] done (multicast commit)
    uint32_t* s_tmem = (uint32_t*)(bars + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int half = N / 2;
    if (threadIdx.x == 0) {
        ptx::mbar_init(smem_u32(&bars[0]), 8);
        ptx::mbar_init(smem_u32(&bars[1]), 1);
        ptx::fence_barrier_init();
    }
    if (warp == 0) ptx::tmem_alloc_pair(smem_u32(s_tmem), 256);
    ptx::tc_fence_before_sync();
    ptx::cluster_sync();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = *s_tmem;
    fill_swizzled(sA, A + (size_t)rank * 128 * KC, 128, threadIdx.x, 128);
    fill_swizzled(sB, B + (size_t)rank * half * KC, half, threadIdx.x, 128);
    ptx::fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive_cluster(ptx::map_to_cta(smem_u32(&bars[0]), 0));
    if (rank == 0 && threadIdx.x == 0) {
        ptx::mbar_wait_cluster(smem_u32(&bars[0]), 0);
        ptx::tc_fence_after_sync();
        const uint32_t idesc = ptx::umma_idesc_bf16(256, N);
        const uint64_t ad = ptx::umma_desc_sw128(smem_u32(sA)), bd = ptx::umma_desc_sw128(smem_u32(sB));
        for (int ks = 0; ks < 4; ++ks) ptx::umma_bf16_pair(tmem, ad + 2 * ks, bd + 2 * ks, idesc, ks > 0);
        ptx::umma_commit_pair(smem_u32(&bars[1]), 3);
    }
    ptx::mbar_wait(smem_u32(&bars[1]), 0);
    ptx::tc_fence_after_sync();
    const int row = rank * 128 + warp * 32 + lane;
    for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v[16];
        ptx::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
        ptx::tmem_ld_wait();
        for (int c = 0; c < 16; ++c) D[(size_t)row * N + c0 + c] = __uint_as_float(v[c]);
    }
    ptx::tc_fence_before_sync();
    ptx::cluster_sync();
    if (warp == 0) ptx::tmem_dealloc_pair(tmem, 256);
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 2; } } while (0)

int main(int argc, char** argv) {
    const char* mode = argc > 1 ? argv[1] : "one";
    const int N = argc > 2 ? atoi(argv[2]) : 256;
    const int M = strcmp(mode, "pair") == 0 ? 256 : 128;
    std::vector<__nv_bfloat16> hA((size_t)M * KC), hB((size_t)N * KC);
    std::vector<float> fA(hA.size()), fB(hB.size());
    srand(7);
    for (size_t i = 0; i < hA.size(); ++i) { hA[i] = __float2bfloat16((rand() % 2001 - 1000) / 1000.f); fA[i] = __bfloat162float(hA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { hB[i] = __float2bfloat16((rand() % 2001 - 1000) / 1000.f); fB[i] = __bfloat162float(hB[i]); }
    __nv_bfloat16 *dA, *dB; float* dD;
    CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dB, hB.size() * 2)); CK(cudaMalloc(&dD, (size_t)M * N * 4));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0xff, (size_t)M * N * 4));
    const int smem = 16384 + 256 * 128 + 64 + 1024;
    if (M == 128) {
        CK(cudaFuncSetAttribute(one_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        one_kernel<<<1, 128, smem>>>(dA, dB, dD, N);
    } else {
        CK(cudaFuncSetAttribute(pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        pair_kernel<<<2, 128, smem>>>(dA, dB, dD, N);
    }
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<float> hD((size_t)M * N);
    CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0, maxerr_swapped = 0;
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            double acc = 0, acc2 = 0;
            int n2 = (n + N / 2) % N;
            for (int k = 0; k < KC; ++k) { acc += (double)fA[m * KC + k] * fB[n * KC + k]; acc2 += (double)fA[m * KC + k] * fB[n2 * KC + k]; }
            double e = fabs(acc - hD[(size_t)m * N + n]), e2 = fabs(acc2 - hD[(size_t)m * N + n]);
            if (e > maxerr) maxerr = e;
            if (e2 > maxerr_swapped) maxerr_swapped = e2;
        }
    printf("umma_selftest %s N=%d: max |err| = %.3e (B halves swapped hypothesis: %.3e)  D[0][0]=%f D[last]=%f -> %s\n", mode, N,
           maxerr, maxerr_swapped, hD[0], hD[hD.size() - 1], maxerr < 1e-3 ? "PASS" : "FAIL");
    return maxerr < 1e-3 ? 0 : 1;
}
