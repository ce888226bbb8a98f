// Issue-slot microbenchmark: how do packed FP32 (FFMA2), scalar FP32, ALU, IMAD and LDS instructions share one
// scheduler on sm_100a?  Every variant runs ITER iterations of a body with NF floating-point instructions over
// independent accumulators, optionally interleaved with NI "other" instructions; cycles per warp-iteration per
// scheduler follow from clock64() with 8 warps per scheduler resident.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITER 4096

template <int FP, int OTHER, int NOTH>   // FP: 0 none, 1 FFMA2, 2 FFMA, 3 FMUL2, 4 FADD2, 5 FMUL, 6 FADD;  OTHER: 0 none 1 LOP3 2 IMAD 3 LDS 4 IADD3 5 SHF
__global__ void __launch_bounds__(256, 4) k(float *out, long long *cyc, float a, float b, int ki) {
    __shared__ float sm[1024];
    sm[threadIdx.x] = a; sm[threadIdx.x + 256] = b; sm[threadIdx.x + 512] = a; sm[threadIdx.x + 768] = b;
    __syncthreads();
    unsigned long long acc[8];
    float f[8];
    unsigned x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i] = ((unsigned long long)__float_as_uint(a + i) << 32) | __float_as_uint(b + i); f[i] = a + i; x[i] = threadIdx.x + i; }
    unsigned long long ab = ((unsigned long long)__float_as_uint(a) << 32) | __float_as_uint(a);
    unsigned long long bb = ((unsigned long long)__float_as_uint(b) << 32) | __float_as_uint(b);
    unsigned saddr = (unsigned)__cvta_generic_to_shared(sm) + threadIdx.x * 4;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (FP == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(ab), "l"(bb));
                if (FP == 2) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(a), "f"(b));
                if (FP == 3) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(acc[i]) : "l"(ab));
                if (FP == 4) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(acc[i]) : "l"(bb));
                if (FP == 5) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(a));
                if (FP == 6) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(b));
                if (i < NOTH) {
                    if (OTHER == 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(ki), "r"(it));
                    if (OTHER == 2) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(ki), "r"(it));
                    if (OTHER == 3) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(saddr + ((i & 3) << 10)) : "memory"); x[i] ^= __float_as_uint(v); }
                    if (OTHER == 4) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(ki));
                    if (OTHER == 5) asm volatile("shf.l.wrap.b32 %0, %0, %0, %1;" : "+r"(x[i]) : "r"(ki));
                    if (OTHER == 6) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(a));
                    if (OTHER == 7) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(b));
                    if (OTHER == 8) asm volatile("{ .reg .b64 t; mov.b64 t, {%0, %0}; mul.rn.f32x2 t, t, %1; mov.b64 {%0, _}, t; }" : "+f"(f[i]) : "l"(ab));
                }
            }
        }
    }
    long long t1 = clock64();
    float s = 0; unsigned y = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { s += __uint_as_float((unsigned)acc[i]) + __uint_as_float((unsigned)(acc[i] >> 32)) + f[i]; y ^= x[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)y;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int FP, int OTHER, int NOTH>
void run(const char *name, float *out, long long *cyc) {
    int grid = 148 * 4;            // 4 CTAs x 8 warps per SM = 8 warps per scheduler
    k<FP, OTHER, NOTH><<<grid, 256>>>(out, cyc, 1.0f, 0.0f, 3);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<FP, OTHER, NOTH><<<grid, 256>>>(out, cyc, 1.0f, 0.0f, 3);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[148 * 4]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < grid; ++i) mean += h[i]; mean /= grid;
    // per scheduler: 8 warps; cycles per (warp-iteration) = mean / ITER / 8
    int nfp = FP ? 16 : 0, noth = OTHER ? 2 * NOTH : 0;
    printf("%-34s fp %2d other %2d per warp-iter: %.2f cycles per warp-iter per scheduler  (%.3f ms)\n", name, nfp, noth, mean / ITER / 8.0, ms);
}

int main() {
    float *out; long long *cyc;
    cudaMalloc(&out, 148 * 4 * 256 * 4); cudaMalloc(&cyc, 148 * 4 * 8);
    run<1, 0, 0>("FFMA2", out, cyc);
    run<2, 0, 0>("FFMA", out, cyc);
    run<3, 0, 0>("FMUL2", out, cyc);
    run<4, 0, 0>("FADD2", out, cyc);
    run<5, 0, 0>("FMUL", out, cyc);
    run<6, 0, 0>("FADD", out, cyc);
    run<0, 1, 8>("LOP3 x16", out, cyc);
    run<0, 2, 8>("IMAD x16", out, cyc);
    run<0, 3, 8>("LDS x16", out, cyc);
    run<0, 4, 8>("IADD x16", out, cyc);
    run<0, 5, 8>("SHF x16", out, cyc);
    run<1, 1, 4>("FFMA2 + LOP3 x8", out, cyc);
    run<1, 1, 8>("FFMA2 + LOP3 x16", out, cyc);
    run<1, 2, 4>("FFMA2 + IMAD x8", out, cyc);
    run<1, 2, 8>("FFMA2 + IMAD x16", out, cyc);
    run<1, 3, 4>("FFMA2 + LDS x8", out, cyc);
    run<1, 3, 8>("FFMA2 + LDS x16", out, cyc);
    run<1, 4, 8>("FFMA2 + IADD x16", out, cyc);
    run<1, 5, 8>("FFMA2 + SHF x16", out, cyc);
    run<1, 6, 2>("FFMA2 x16 + FMUL x4", out, cyc);
    run<1, 6, 4>("FFMA2 x16 + FMUL x8", out, cyc);
    run<1, 6, 8>("FFMA2 x16 + FMUL x16", out, cyc);
    run<1, 7, 8>("FFMA2 x16 + FADD x16", out, cyc);
    run<3, 7, 8>("FMUL2 x16 + FADD x16", out, cyc);
    run<1, 8, 8>("FFMA2 x16 + FMUL2(half used) x16", out, cyc);
    run<5, 7, 8>("FMUL x16 + FADD x16", out, cyc);
    run<2, 1, 8>("FFMA + LOP3 x16", out, cyc);
    run<2, 2, 8>("FFMA + IMAD x16", out, cyc);
    run<3, 1, 8>("FMUL2 + LOP3 x16", out, cyc);
    run<5, 1, 8>("FMUL + LOP3 x16", out, cyc);
    run<6, 1, 8>("FADD + LOP3 x16", out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
