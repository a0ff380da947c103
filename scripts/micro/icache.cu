// Microbenchmark: cost of executing cold straight-line code once (instruction fetch) on one warp.
#include <cstdio>
#include <cuda_runtime.h>
#define R4(x) x x x x
#define R16(x) R4(R4(x))
#define R256(x) R16(R16(x))
#define R1024(x) R4(R256(x))
#define BODY a = fmaf(a, c, d); b = fmaf(b, c, d); e = fmaf(e, c, d); f = fmaf(f, c, d);
template <int REP>
__global__ void k(float* out, long long* t, float c, float d)
{
    float a = out[0], b = out[1], e = out[2], f = out[3];
    long long t0 = clock64();
    if (REP >= 1) { R1024(BODY) }
    if (REP >= 2) { R1024(BODY) }
    if (REP >= 4) { R1024(BODY) R1024(BODY) }
    if (REP >= 8) { R1024(BODY) R1024(BODY) R1024(BODY) R1024(BODY) }
    long long t1 = clock64();
    out[4 + threadIdx.x] = a + b + e + f;
    if (threadIdx.x == 0) t[blockIdx.x] = t1 - t0;
}
template <int REP> void run(float* out, long long* t, const char* name)
{
    for (int it = 0; it < 3; ++it) {
        k<REP><<<1, 32>>>(out, t, 1.0001f, 0.5f);
        cudaDeviceSynchronize();
        long long h; cudaMemcpy(&h, t, 8, cudaMemcpyDeviceToHost);
        printf("%s launch %d: %d FFMA, %lld cycles, %.2f cyc/instr\n", name, it, REP * 4096, h, (double)h / (REP * 4096));
    }
}
int main()
{
    float* out; long long* t;
    cudaMalloc(&out, 4096); cudaMalloc(&t, 4096); cudaMemset(out, 0, 4096);
    run<1>(out, t, "4K instr (64 KB)");
    run<2>(out, t, "8K instr (128 KB)");
    run<8>(out, t, "32K instr (512 KB)");
    return 0;
}
