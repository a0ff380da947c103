// Helpers of the critical-path schedule (factor_cp in gauss_cuda.cu).
//
// The kb x kb diagonal region A11 of an outer block is factored inside an EXTENDED scratch matrix
//
//        E  =  [ A11  I ]        after eliminating the first kb pivots        [ L11|U11   inv(L11) ]
//              [ I    0 ]        (the unchanged in-panel chain of kernels)    [ inv(U11)  -inv(A11) ]
//
// so the same diag-factor / U-row / L-col / rank-128 kernels that factor A11 (getPivot + computeGauss restricted to
// the block, gauss_internal_input.c:100-111, :129-136) also deliver both triangular inverses: the U-row solve applied
// to the identity columns is L11^-1, the L-col solve applied to the identity rows is U11^-1.  With them the two panel
// solves of the outer block become single GEMMs on the tile engine (L21 = A21 inv(U11), U12 = inv(L11) A12) instead
// of 2 x kb/128 dependent launches, which is what shortens the multi-GPU critical path.
#include "gauss_internal.cuh"

namespace gb200 {

__device__ __forceinline__ void cp_split(float x, float& hi, float& lo)
{
    uint32_t h, l;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    hi = __uint_as_float(h);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(x - hi));
    lo = __uint_as_float(l);
}

// E[0:2kb, 0:2kb] (leading dimension lde) <- [[A11, I], [I, 0]], A11 = the kb x kb block at `a11` (ld).
__global__ void __launch_bounds__(256)
ext_prepare_kernel(const float* __restrict__ a11, int ld, float* __restrict__ E, int lde, int kb)
{
    const int c4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int r = blockIdx.y;
    if (c4 >= 2 * kb) return;
    float4 v;
    if (r < kb && c4 < kb) {
        v = *reinterpret_cast<const float4*>(a11 + (size_t)r * ld + c4);
    } else {
        const int d = (r < kb) ? (r + kb) : (c4 < kb ? r - kb : -8);      // column of the 1 in this row's identity part
        v.x = (c4 == d) ? 1.f : 0.f;
        v.y = (c4 + 1 == d) ? 1.f : 0.f;
        v.z = (c4 + 2 == d) ? 1.f : 0.f;
        v.w = (c4 + 3 == d) ? 1.f : 0.f;
    }
    *reinterpret_cast<float4*>(E + (size_t)r * lde + c4) = v;
}

void launch_ext_prepare(const float* a11, int ld, float* E, int lde, int kb, cudaStream_t s)
{
    dim3 grid((2 * kb / 4 + 255) / 256, 2 * kb);
    ext_prepare_kernel<<<grid, 256, 0, s>>>(a11, ld, E, lde, kb);
    ++g_launch_count;
}

// After the chain: one launch that
//   (y-blocks [0, kb/32))        writes L11|U11 back into the matrix, and inv(L11) = E[0:kb, kb:2kb] as a row-major
//                                hi/lo split (il_*: the A operand of U12 = inv(L11) A12 and the B operand of its transpose)
//   (y-blocks [kb/32, 2kb/32))   writes inv(U11) = E[kb:2kb, 0:kb] TRANSPOSED as a hi/lo split (iut_*[n*ldo + k] =
//                                inv(U11)[k][n]: the K-major B operand of L21 = A21 inv(U11))
// 32 x 32 tiles, 32 x 8 threads.
__global__ void __launch_bounds__(256)
ext_finish_kernel(const float* __restrict__ E, int lde, int kb, float* __restrict__ a11, int ld,
                  float* __restrict__ il_hi, float* __restrict__ il_lo, float* __restrict__ iut_hi, float* __restrict__ iut_lo, int ldo)
{
    __shared__ float t[32][33];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int c0 = blockIdx.x * 32;
    const int nby = kb / 32;
    if ((int)blockIdx.y < nby) {
        const int r0 = blockIdx.y * 32;
        for (int i = ty; i < 32; i += 8) {
            const int r = r0 + i, c = c0 + tx;
            a11[(size_t)r * ld + c] = E[(size_t)r * lde + c];
            float h, l;
            cp_split(E[(size_t)r * lde + kb + c], h, l);
            il_hi[(size_t)r * ldo + c] = h;
            il_lo[(size_t)r * ldo + c] = l;
        }
    } else {
        const int r0 = (blockIdx.y - nby) * 32;                  // row block of inv(U11)
        for (int i = ty; i < 32; i += 8) t[i][tx] = E[(size_t)(kb + r0 + i) * lde + c0 + tx];
        __syncthreads();
        for (int i = ty; i < 32; i += 8) {
            float h, l;
            cp_split(t[tx][i], h, l);                            // element (row r0+tx, column c0+i) -> out[c0+i][r0+tx]
            iut_hi[(size_t)(c0 + i) * ldo + r0 + tx] = h;
            iut_lo[(size_t)(c0 + i) * ldo + r0 + tx] = l;
        }
    }
}

void launch_ext_finish(const float* E, int lde, int kb, float* a11, int ld,
                       float* il_hi, float* il_lo, float* iut_hi, float* iut_lo, int ldo, cudaStream_t s)
{
    dim3 grid(kb / 32, 2 * (kb / 32));
    ext_finish_kernel<<<grid, dim3(32, 8), 0, s>>>(E, lde, kb, a11, ld, il_hi, il_lo, iut_hi, iut_lo, ldo);
    ++g_launch_count;
}

} // namespace gb200
