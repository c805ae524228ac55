// Micro-benchmark: ex2.approx.ftz.f32 (MUFU.EX2) and FFMA issue rates per SM per clock on this GPU.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu mufu.cu && ./mufu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float *out, int iters, long long *clk)
{
    float a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 1e-3f + i * 0.1f;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            else if (MODE == 1) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i]));
            else { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])); asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i])); asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i])); asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i])); }
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, int warps_per_sm)
{
    int sms = 148, threads = warps_per_sm * 32, iters = 4096;
    float *out; long long *clk;
    cudaMalloc(&out, sms * threads * 4); cudaMalloc(&clk, sms * 8);
    k<MODE><<<sms, threads>>>(out, 16, clk);
    k<MODE><<<sms, threads>>>(out, iters, clk);
    long long h[148]; cudaMemcpy(h, clk, sms * 8, cudaMemcpyDeviceToHost);
    double c = 0; for (int i = 0; i < sms; i++) c += h[i]; c /= sms;
    double ops = (double)iters * 8 * threads * (MODE == 2 ? 1 : 1);
    printf("%-28s warps/SM %2d: %.1f clk, %.2f %s/clk/SM\n", name, warps_per_sm, c, ops / c, MODE == 1 ? "fma" : "ex2");
    cudaFree(out); cudaFree(clk);
}

int main()
{
    for (int w : {4, 8, 16, 32}) run<0>("ex2 only", w);
    for (int w : {4, 8, 16}) run<1>("ffma only", w);
    for (int w : {4, 8, 16}) run<2>("ex2 + 3 ffma", w);
    return 0;
}
