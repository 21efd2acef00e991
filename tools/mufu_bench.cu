// Micro-benchmark: MUFU.EX2 / MUFU.RCP / polynomial-exp2 (FMA pipe) throughput per SM on the box's GPU.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_bench mufu_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float *out, int iters, long long *cyc) {
    float v[8];
    for (int i = 0; i < 8; ++i) v[i] = 0.001f * (threadIdx.x + i);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
            if (MODE == 1) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
            if (MODE == 2) {   // degree-5 polynomial exp2 on the FMA pipe (Cody-Waite): round, poly, exponent add
                float x = v[i];
                float n = rintf(x);
                float f = x - n;
                float p = 1.3333558e-3f;
                p = fmaf(p, f, 9.6181291e-3f); p = fmaf(p, f, 5.5504109e-2f); p = fmaf(p, f, 2.4022651e-1f);
                p = fmaf(p, f, 6.9314718e-1f); p = fmaf(p, f, 1.0f);
                v[i] = __int_as_float(__float_as_int(p) + ((int)n << 23)) * 0.25f;
            }
            if (MODE == 3) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i])); v[i] = fmaf(v[i], 0.5f, 0.1f); v[i] = fmaf(v[i], 0.5f, 0.1f); v[i] = fmaf(v[i], 0.5f, 0.1f); v[i] = fmaf(v[i], 0.5f, 0.1f); v[i] = fmaf(v[i], 0.5f, 0.1f); v[i] = fmaf(v[i], 0.5f, 0.1f); }
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < 8; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    float *out; long long *cyc, h;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 4096;
    const char *names[4] = {"ex2", "rcp", "poly-exp2 (FMA pipe)", "ex2 + 6 FMA"};
    for (int warps = 4; warps <= 16; warps *= 2)
        for (int mode = 0; mode < 4; ++mode) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) k<0><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 1) k<1><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 2) k<2><<<148, warps * 32>>>(out, iters, cyc);
                if (mode == 3) k<3><<<148, warps * 32>>>(out, iters, cyc);
                cudaDeviceSynchronize();
            }
            cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
            double ops = (double)iters * 8 * warps * 32;
            printf("%-22s warps/SM %2d: %.2f results/clk/SM (%lld clk)\n", names[mode], warps, ops / h, h);
        }
    return 0;
}
