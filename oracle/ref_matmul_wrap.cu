/*
 * oracle/ref_matmul_wrap.cu -- TEST INFRASTRUCTURE ONLY.
 *
 * Gives the reference's own matmul code a C ABI so tests and bench.py can run
 * it unmodified: the translation unit below is the reference file itself,
 * pulled in by path at build time (nothing is copied into this repo), with its
 * main() renamed.  Built by oracle/Makefile into oracle/_ref/libmatmul_ref.so
 * only where /root/reference exists.
 *
 *   ref_seq_matmul  -> seq_matmul   CUDA_and_OpenMP/Version-2/cuda_matmul.cu:28-39
 *   ref_omp_matmul  -> omp_matmul   cuda_matmul.cu:42-57
 *   ref_gpu_matmul  -> gpu_matmul   cuda_matmul.cu:89-101, launched with the
 *                                   reference's own grid/block (cuda_matmul.cu:155)
 */
#include <cuda_runtime.h>
#include <math.h>

void seq_matmul(const float* A, const float* B, float* C, int nsize);
void omp_matmul(const float* A, const float* B, float* C, int nsize);
__global__ void gpu_matmul(float* A, float* B, float* C, int nsize);

extern "C" void ref_seq_matmul(const float* A, const float* B, float* C, int n) { seq_matmul(A, B, C, n); }
extern "C" void ref_omp_matmul(const float* A, const float* B, float* C, int n) { omp_matmul(A, B, C, n); }

/* Device pointers in, kernel only (no malloc/copies): returns the kernel's
 * CUDA-event time in milliseconds, or a negative cudaError_t. */
extern "C" float ref_gpu_matmul(float* dA, float* dB, float* dC, int n)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    dim3 grid(n, (unsigned)ceil(n / 1024.0));
    int block = n < 1024 ? n : 1024;
    cudaEventRecord(e0);
    gpu_matmul<<<grid, block>>>(dA, dB, dC, n);
    cudaEventRecord(e1);
    cudaError_t rc = cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (rc != cudaSuccess) return -(float)rc;
    rc = cudaGetLastError();
    return rc == cudaSuccess ? ms : -(float)rc;
}

#define main ref_matmul_main
#include REF_MATMUL_SOURCE
