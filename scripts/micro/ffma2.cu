// Microbenchmark: scalar FFMA vs packed FFMA2 (fma.rn.f32x2) throughput on sm_100a, all SMs busy.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float c0, float c1)
{
    float2 a[8], m = make_float2(c0, c0), d = make_float2(c1, c1);
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = make_float2(threadIdx.x * 1e-3f + i, blockIdx.x * 1e-3f - i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) { a[i].x = fmaf(a[i].x, m.x, d.x); a[i].y = fmaf(a[i].y, m.y, d.y); }
            else           { a[i] = __ffma2_rn(a[i], m, d); }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, float* out)
{
    const int iters = 20000, blocks = 148 * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(out, 100, 1.0001f, 0.5f);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(out, iters, 1.0001f, 0.5f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double fma = (double)blocks * 256 * iters * 16;
    printf("%s: %.3f ms, %.1f TFLOP/s (2 flop per FMA)\n", name, ms, 2.0 * fma / (ms * 1e-3) * 1e-12);
}
int main()
{
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<0>("scalar FFMA ", out);
    run<1>("packed FFMA2", out);
    return 0;
}
