// Micro-benchmark: per-SM throughput of MUFU.SIN/COS, MUFU.EX2, F2FP.BF16 pack and FFMA (ops / clk / SM).
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
    float a[8];
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 0.001f + i;
    unsigned acc = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) a[i] = __sinf(a[i]);
            if (MODE == 1) a[i] = __cosf(a[i]) + __sinf(a[i]);
            if (MODE == 2) a[i] = exp2f(a[i]);
            if (MODE == 3) { unsigned r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(a[i]), "f"(a[(i + 1) & 7])); acc ^= r; }
            if (MODE == 4) a[i] = fmaf(a[i], 1.0001f, 0.5f);
        }
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
    const char* names[] = {"sin (FMUL+MUFU.SIN)", "sin+cos (2 MUFU)", "ex2", "F2FP.BF16 pack", "FFMA"};
    const int iters = 2000;
    for (int warps = 8; warps <= 32; warps *= 2)
        for (int mode = 0; mode < 5; ++mode) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) k<0><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 1) k<1><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 2) k<2><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 3) k<3><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 4) k<4><<<148, warps * 32>>>(out, iters, cyc);
                cudaDeviceSynchronize();
            }
            double ops = (double)iters * 8 * warps * 32 * (mode == 1 ? 2 : 1);
            printf("warps/SM=%2d %-22s %7.2f ops/clk/SM\n", warps, names[mode], ops / (double)*cyc);
        }
    return 0;
}
