// Micro-benchmark: per-SM throughput of the legacy tensor path (mma.sync.m16n8k16 f16, SASS HMMA.16816) on sm_100a,
// with fp32 and fp16 accumulators, alone and mixed with ldmatrix.x4 (one per 4 MMAs: the K2c inner loop's ratio).
// Prints MMA / clk / SM and the equivalent dense TFLOP/s at the SM clock the run saw.
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

__device__ __forceinline__ void mma_f32(float (&d)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void mma_f16(unsigned (&d)[2], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f16.f16.f16.f16 {%0,%1}, {%2,%3,%4,%5}, {%6,%7}, {%0,%1};"
                 : "+r"(d[0]), "+r"(d[1])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void ldsm4(unsigned (&r)[4], unsigned addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}

// MODE 0: f32 acc, 16 independent accumulators; 1: f16 acc; 2: f32 acc + one ldmatrix.x4 per 4 MMA; 3: f16 acc + ldmatrix
template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
    __shared__ __align__(16) unsigned char sm[16 * 1024];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) ((unsigned*)sm)[i] = 0x3c003c00u;
    __syncthreads();
    unsigned a[4][4], b[4][2];
    for (int i = 0; i < 4; ++i) {
        for (int j = 0; j < 4; ++j) a[i][j] = 0x3c003c00u + threadIdx.x + i + j;
        for (int j = 0; j < 2; ++j) b[i][j] = 0x38003800u + i + j;
    }
    float d32[16][4] = {};
    unsigned d16[16][2] = {};
    const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + (threadIdx.x & 31) * 16 + ((threadIdx.x >> 5) & 7) * 2048;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            if (MODE >= 2) ldsm4(a[s], base + s * 512);
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                if (MODE == 0 || MODE == 2) mma_f32(d32[s * 4 + t], a[s], b[t]);
                else mma_f16(d16[s * 4 + t], a[s], b[t]);
            }
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < 16; ++i) s += d32[i][0] + d32[i][3] + (float)d16[i][0] + (float)d16[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const char* names[] = {"f32 acc", "f16 acc", "f32 acc + ldmatrix.x4/4", "f16 acc + ldmatrix.x4/4"};
    const int iters = 2000;
    for (int warps = 4; warps <= 32; warps *= 2)
        for (int mode = 0; mode < 4; ++mode) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) k<0><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 1) k<1><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 2) k<2><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 3) k<3><<<148, warps * 32>>>(out, iters, cyc);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            const double mmas = (double)iters * 16 * warps, per_clk = mmas / (double)*cyc;
            printf("warps/SM=%2d %-26s %6.3f HMMA.16816/clk/SM = %6.0f dense FMA/clk/SM = %6.0f TFLOP/s at %d MHz\n", warps,
                   names[mode], per_clk, per_clk * 2048, per_clk * 4096 * 148 * clk_khz * 1e3 / 1e12, clk_khz / 1000);
        }
    return 0;
}
