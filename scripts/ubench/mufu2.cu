// Micro-benchmark: candidate instruction mixes of the attention kernel's exponential phase, in "score pairs per clk per SM".
//   mode 0  current:  2 x ex2.approx.ftz.f32 + 1 x cvt.rn.f16x2.f32 (pack)                     per pair
//   mode 1  packed:   1 x cvt.rn.f16x2.f32 + 1 x ex2.approx.f16x2                          per pair
//   mode 2  ex2.f16x2 alone        mode 3  cvt.rn.f16x2.f32 alone      mode 4  ex2.f32 alone (x2 per pair)
//   mode 5  packed + row sum on the mixed-precision FMA (2 x fma.rn.f32.f16 per pair) + 2 x ffma for the scaling
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu2 mufu2.cu && ./mufu2
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

template <int MODE>
__global__ void k(float *out, int iters, long long *clk)
{
    float a[8], b[8], s0 = 0.f, s1 = 0.f;
    unsigned h[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = -(threadIdx.x * 1e-3f + i * 0.1f); b[i] = a[i] - 0.5f; h[i] = 0x3c003c00u; }
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) {
                float p0, p1;
                asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(p0) : "f"(a[i]));
                asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(p1) : "f"(b[i]));
                asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(p1), "f"(p0));
                a[i] -= 1e-6f;
            } else if (MODE == 1) {
                unsigned x;
                asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(x) : "f"(b[i]), "f"(a[i]));
                asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(h[i]) : "r"(x));
                a[i] -= 1e-6f;
            } else if (MODE == 2) {
                asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
            } else if (MODE == 3) {
                asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(b[i]), "f"(a[i]));
                a[i] -= 1e-6f;
            } else if (MODE == 4) {
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b[i]));
            } else {
                unsigned x;
                float x0 = fmaf(a[i], 1.0001f, -0.25f), x1 = fmaf(b[i], 1.0001f, -0.25f);
                asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(x) : "f"(x1), "f"(x0));
                asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(h[i]) : "r"(x));
                asm volatile("{\n\t.reg .b16 lo, hi, one;\n\tmov.b32 {lo, hi}, %2;\n\tmov.b16 one, 0x3c00;\n\t"
                             "fma.rn.f32.f16 %0, lo, one, %0;\n\tfma.rn.f32.f16 %1, hi, one, %1;\n\t}" : "+f"(s0), "+f"(s1) : "r"(h[i]));
                a[i] -= 1e-6f;
            }
        }
    }
    long long t1 = clock64();
    float s = s0 + s1;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i] + b[i] + __half2float(__ushort_as_half((unsigned short)(h[i] & 0xffff)));
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
    double pairs = (double)iters * 8 * threads;
    printf("%-44s warps/SM %2d: %.2f pairs/clk/SM\n", name, warps_per_sm, pairs / c);
    cudaFree(out); cudaFree(clk);
}

int main()
{
    for (int w : {8, 16}) run<0>("current: 2 ex2.f32 + pack", w);
    for (int w : {8, 16}) run<1>("packed: pack + ex2.f16x2", w);
    for (int w : {8, 16}) run<2>("ex2.f16x2 alone", w);
    for (int w : {8, 16}) run<3>("cvt.rn.f16x2.f32 alone", w);
    for (int w : {8, 16}) run<4>("2 x ex2.f32 alone", w);
    for (int w : {8, 16}) run<5>("packed + 2 ffma + 2 fhfma (full phase)", w);
    return 0;
}
