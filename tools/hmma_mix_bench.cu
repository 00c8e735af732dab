// Micro-benchmark: the consumer loop of K3r (csrc/k2_gse_embed_cheb.cu) rebuilt stage by stage, to see which part of
// the instruction mix costs tensor-pipe time.  One "m-tile" per iteration = 12 f16-accumulate HMMA.16816 (3 A fragments x
// 4 B fragments) [+ 3 ldmatrix.x4] [+ 8 three-input half2 max] [+ 4 f32-accumulate HMMA, two dependent chains of two]
// [+ 4 FADD + 4 STS].  Prints clocks per HMMA and SM sub-partition (8.0 = the pipe's peak, tools/hmma_bench.cu).
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

__device__ __forceinline__ void mma_h_new(unsigned (&d)[2], const unsigned (&a)[4], const unsigned (&b)[2]) {
    const unsigned z = 0u;
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f16.f16.f16.f16 {%0,%1}, {%2,%3,%4,%5}, {%6,%7}, {%8,%8};"
                 : "=r"(d[0]), "=r"(d[1]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "r"(z));
}
__device__ __forceinline__ void mma_h_acc(unsigned (&d)[2], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f16.f16.f16.f16 {%0,%1}, {%2,%3,%4,%5}, {%6,%7}, {%0,%1};"
                 : "+r"(d[0]), "+r"(d[1]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void mma_new(float (&d)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
    const float z = 0.f;
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(z));
}
__device__ __forceinline__ void mma_acc(float (&d)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm4(unsigned (&r)[4], unsigned addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ unsigned hmax2_u(unsigned x, unsigned y) {
    unsigned r;
    asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(y));
    return r;
}

enum { M_ACC = 0, M_ACC_LDSM, M_NEW_MAX, M_NEW_MAX_LDSM, M_SCORE, M_SCORE_INDEP, M_SCORE_STS, M_SCORE_HI_ONLY, M_PIPELINED, M_COUNT };
static const char* kNames[M_COUNT] = {
    "12 HMMA.F16 accumulate in place",
    "  + 3 ldmatrix.x4",
    "12 HMMA.F16 (C = 0) + 8 max3.f16x2",
    "  + 3 ldmatrix.x4",
    "  + 4 HMMA.F32 (2 chains of 2) + 4 FADD",
    "  same, 4 independent HMMA.F32 + 12 FADD",
    "  + 4 STS (the shipped m-tile)",
    "  hi half only: 2 HMMA.F32 + 4 FADD + 4 STS",
    "  shipped m-tile, software-pipelined by one",
};

template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
    __shared__ __align__(16) unsigned char sm[32 * 1024];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) ((unsigned*)sm)[i] = 0x3c003c00u;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned a[3][4], b[4][2], qh[4], ql[4];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 4; ++j) a[i][j] = 0x3c003c00u + threadIdx.x + i + j;
    for (int i = 0; i < 4; ++i) {
        for (int j = 0; j < 2; ++j) b[i][j] = 0x38003800u + i + j;
        qh[i] = 0x34003400u + i;
        ql[i] = 0x14001400u + i;
    }
    unsigned c[3][4][2] = {};
    float keep[4] = {};
    const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + lane * 16 + (warp & 7) * 2048;
    float* const st = reinterpret_cast<float*>(sm + 16 * 1024) + (warp & 15) * 256 + lane;
    constexpr bool LDSM = MODE == M_ACC_LDSM || MODE >= M_NEW_MAX_LDSM;
    auto expand_new = [&]() {
#pragma unroll
        for (int s = 0; s < 3; ++s)
#pragma unroll
            for (int t = 0; t < 4; ++t) mma_h_new(c[s][t], a[s], b[t]);
    };
    auto fold = [&](unsigned (&e0)[4], unsigned (&e1)[4]) {
#pragma unroll
        for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int r = 0; r < 2; ++r) (t < 2 ? e0 : e1)[(t & 1) * 2 + r] = hmax2_u(hmax2_u(c[0][t][r], c[1][t][r]), c[2][t][r]);
    };
    long long t0 = clock64();
    if (MODE == M_PIPELINED) {
        float scp[4] = {}, sc1p[4] = {};
#pragma unroll
        for (int s = 0; s < 3; ++s) ldsm4(a[s], base + s * 512);
        expand_new();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int s = 0; s < 3; ++s) ldsm4(a[s], base + s * 512 + (it & 1) * 16);
            unsigned e0[4], e1[4];
            fold(e0, e1);
            expand_new();
            st[0] = scp[0] + sc1p[0]; st[32] = scp[1] + sc1p[1]; st[64] = scp[2] + sc1p[2]; st[96] = scp[3] + sc1p[3];
            mma_new(scp, e0, qh[0], qh[1]);
            mma_new(sc1p, e1, qh[2], qh[3]);
            mma_acc(scp, e0, ql[0], ql[1]);
            mma_acc(sc1p, e1, ql[2], ql[3]);
        }
        keep[0] = scp[0] + sc1p[1];
    } else {
        for (int it = 0; it < iters; ++it) {
            if (LDSM) {
#pragma unroll
                for (int s = 0; s < 3; ++s) ldsm4(a[s], base + s * 512 + (it & 1) * 16);
            }
            if (MODE <= M_ACC_LDSM) {
#pragma unroll
                for (int s = 0; s < 3; ++s)
#pragma unroll
                    for (int t = 0; t < 4; ++t) mma_h_acc(c[s][t], a[s], b[t]);
                continue;
            }
            expand_new();
            unsigned e0[4], e1[4];
            fold(e0, e1);
            if (MODE <= M_NEW_MAX_LDSM) {  // consume the maxima cheaply: they become the next iteration's B fragments
                b[0][0] = e0[0]; b[1][0] = e0[2]; b[2][0] = e1[0]; b[3][0] = e1[2];
                b[0][1] = e0[1] ^ e0[3]; b[1][1] = e1[1] ^ e1[3];
                continue;
            }
            float sc[4], sc1[4];
            if (MODE == M_SCORE_INDEP) {
                float sc2[4], sc3[4];
                mma_new(sc, e0, qh[0], qh[1]);
                mma_new(sc1, e1, qh[2], qh[3]);
                mma_new(sc2, e0, ql[0], ql[1]);
                mma_new(sc3, e1, ql[2], ql[3]);
#pragma unroll
                for (int r = 0; r < 4; ++r) keep[r] += (sc[r] + sc1[r]) + (sc2[r] + sc3[r]);
                continue;
            }
            mma_new(sc, e0, qh[0], qh[1]);
            mma_new(sc1, e1, qh[2], qh[3]);
            if (MODE != M_SCORE_HI_ONLY) {
                mma_acc(sc, e0, ql[0], ql[1]);
                mma_acc(sc1, e1, ql[2], ql[3]);
            }
            if (MODE == M_SCORE) {
#pragma unroll
                for (int r = 0; r < 4; ++r) keep[r] += sc[r] + sc1[r];
            } else {
                st[0] = sc[0] + sc1[0]; st[32] = sc[1] + sc1[1]; st[64] = sc[2] + sc1[2]; st[96] = sc[3] + sc1[3];
            }
        }
    }
    long long t1 = clock64();
    float s = keep[0] + keep[1] + keep[2] + keep[3] + st[0];
    for (int i = 0; i < 3; ++i)
        for (int t = 0; t < 4; ++t) s += (float)c[i][t][0] + (float)c[i][t][1];
    s += (float)b[0][0] + (float)b[1][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE>
static void run(float* out, long long* cyc, int iters) {
    static const int hmma_per_iter[M_COUNT] = {12, 12, 12, 12, 16, 16, 16, 14, 16};
    for (int warps = 4; warps <= 16; warps *= 2) {
        for (int rep = 0; rep < 2; ++rep) {
            k<MODE><<<148, warps * 32>>>(out, iters, cyc);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
        }
        const double per_smsp = (double)iters * hmma_per_iter[MODE] * warps / 4.0;
        printf("warps/SM=%2d  %-46s %6.2f clk per HMMA and sub-partition  (%5.1f %% of the 8.0 clk peak)\n", warps, kNames[MODE],
               (double)*cyc / per_smsp, 800.0 * per_smsp / (double)*cyc);
    }
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
    const int iters = 4000;
    run<M_ACC>(out, cyc, iters);
    run<M_ACC_LDSM>(out, cyc, iters);
    run<M_NEW_MAX>(out, cyc, iters);
    run<M_NEW_MAX_LDSM>(out, cyc, iters);
    run<M_SCORE>(out, cyc, iters);
    run<M_SCORE_INDEP>(out, cyc, iters);
    run<M_SCORE_STS>(out, cyc, iters);
    run<M_SCORE_HI_ONLY>(out, cyc, iters);
    run<M_PIPELINED>(out, cyc, iters);
    return 0;
}
