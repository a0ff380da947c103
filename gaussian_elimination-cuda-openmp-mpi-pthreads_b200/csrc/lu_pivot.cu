// Partial-pivoting panel factorisation: the reference's EXTERNAL-input variant
// (OpenMP_and_MPI/gauss_openmp/gauss_external_input.c):
//   getPivot      :123-148  row with the largest |a[i][cur]|, i >= cur (first one wins), swapped in
//   computeGauss  :153-182  pivot row divided by the pivot (:161-169), rows below eliminated (:172-181)
// The 128-column panel of all rows below the diagonal is factored in a transposed scratch copy
// PT[c][i] (column c of the panel contiguous over the rows i), so the per-column pivot search and the
// rank-1 updates are coalesced.  Row swaps are recorded in ipiv and applied to the rest of the matrix
// (RHS column included) afterwards; the trailing update then reuses the no-pivot kernels.
#include "gauss_internal.cuh"

namespace gb200 {

// W[row0 + i][col0 + c] -> PT[c * mp + i]   (i < m, c < 128), 32x32 tiles through shared memory
__global__ void panel_pack_kernel(const float* __restrict__ w, int ld, float* __restrict__ pt, int mp, int m)
{
    __shared__ float t[32][33];
    const int i0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;            // 32 x 8
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + r;
        t[r][tx] = i < m ? w[(size_t)i * ld + c0 + tx] : 0.f;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + tx;
        if (i < m) pt[(size_t)(c0 + r) * mp + i] = t[tx][r];
    }
}

// PT -> W, plus (3xTF32 mode) the hi/lo split of the rows below the diagonal block into la_{hi,lo}[(i-128)*ldla + c]
__global__ void panel_unpack_kernel(float* __restrict__ w, int ld, const float* __restrict__ pt, int mp, int m,
                                    float* __restrict__ la_hi, float* __restrict__ la_lo, int ldla)
{
    __shared__ float t[32][33];
    const int i0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + tx;
        t[r][tx] = i < m ? pt[(size_t)(c0 + r) * mp + i] : 0.f;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + r;
        if (i >= m) continue;
        const float v = t[tx][r];
        w[(size_t)i * ld + c0 + tx] = v;
        if (la_hi != nullptr && i >= kNb) {
            uint32_t hb;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(v));
            const float h = __uint_as_float(hb);
            uint32_t lb;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lb) : "f"(v - h));
            la_hi[(size_t)(i - kNb) * ldla + c0 + tx] = h;
            la_lo[(size_t)(i - kNb) * ldla + c0 + tx] = __uint_as_float(lb);
        }
    }
}

// One pivot step inside the packed panel: find the pivot of column j among rows [j, m), swap rows j and p
// across the 128 panel columns, divide the pivot row (columns > j) by the pivot.  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024)
pivot_step_kernel(float* __restrict__ pt, int mp, int m, int j, int grow0, int* __restrict__ ipiv, int* __restrict__ status)
{
    __shared__ float s_val[32];
    __shared__ int s_idx[32];
    __shared__ int s_piv;
    __shared__ float s_pivval;
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const float* col = pt + (size_t)j * mp;
    float best = -1.f;
    int bi = j;
    for (int i = j + tid; i < m; i += 1024) {
        const float a = fabsf(col[i]);
        if (a > best) { best = a; bi = i; }                // strict >, ascending i: the first maximum wins (:127-131)
    }
    for (int off = 16; off > 0; off >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, off);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
        if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
    }
    if (lane == 0) { s_val[wrp] = best; s_idx[wrp] = bi; }
    __syncthreads();
    if (wrp == 0) {
        best = s_val[lane]; bi = s_idx[lane];
        for (int off = 16; off > 0; off >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, off);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) {
            s_piv = bi;
            ipiv[grow0 + j] = grow0 + bi;
            if (!(best > 0.f) || isinf(best)) atomicCAS(status, 0, grow0 + j + 1);   // "The matrix is singular" (:133-136)
        }
    }
    __syncthreads();
    const int p = s_piv;
    if (tid < kNb) {
        float* cc = pt + (size_t)tid * mp;
        float a = cc[j], b = cc[p];
        if (p != j) { cc[j] = b; cc[p] = a; a = b; }       // row swap inside the panel (:139-146)
        if (tid == j) s_pivval = a;                        // the pivot itself
    }
    __syncthreads();
    const float piv = s_pivval;
    if (tid < kNb && tid > j) {
        float* cc = pt + (size_t)tid * mp;
        cc[j] = __fdiv_rn(cc[j], piv);                     // scale the main row (:161-169)
    }
}

// Rank-1 update of the panel columns right of j for the rows below j:  PT[c][i] -= PT[j][i] * PT[c][j]
__global__ void __launch_bounds__(256)
panel_update_kernel(float* __restrict__ pt, int mp, int m, int j)
{
    __shared__ float u[kNb];
    const int tid = threadIdx.x;
    if (tid < kNb) u[tid] = tid > j ? pt[(size_t)tid * mp + j] : 0.f;
    __syncthreads();
    const int i = j + 1 + blockIdx.x * 256 + tid;
    if (i >= m) return;
    const float l = pt[(size_t)j * mp + i];
    if (l == 0.f) return;                                  // sparse inputs: most multipliers are zero
    for (int c = j + 1; c < kNb; ++c) {
        float* q = pt + (size_t)c * mp + i;
        *q = fmaf(-l, u[c], *q);                           // (:177-179)
    }
}

// Apply the panel's row interchanges to `ncols` columns starting at w (row 0 of w = global row grow0 - roff ...):
// thread per column, swaps applied in pivot order.  rows are addressed globally: w points at W[0][col0].
__global__ void laswp_kernel(float* __restrict__ w, int ld, int ncols, const int* __restrict__ ipiv, int grow0)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncols) return;
    for (int j = 0; j < kNb; ++j) {
        const int r = grow0 + j, p = ipiv[r];
        if (p != r) {
            const float a = w[(size_t)r * ld + c], b = w[(size_t)p * ld + c];
            w[(size_t)r * ld + c] = b;
            w[(size_t)p * ld + c] = a;
        }
    }
}

// Factor the panel W[k0.., k0..k0+128) of an [np x ld] single-GPU system with partial pivoting.
// pt: scratch of 128 * mp floats (mp >= np - k0); on return W holds L (pivots included) and unit-upper U in the
// panel columns, ipiv[k0..k0+128) the global pivot rows, and the interchanges have been applied to the
// columns right of the panel (the RHS column among them).
void launch_panel_pivot(float* W, int ld, int np, int k0, float* pt, int mp, int* ipiv, int* status,
                        float* la_hi, float* la_lo, int ldla, cudaStream_t s)
{
    const int m = np - k0;
    float* wp = W + (size_t)k0 * ld + k0;
    dim3 tgrid((m + 31) / 32, kNb / 32);
    panel_pack_kernel<<<tgrid, dim3(32, 8), 0, s>>>(wp, ld, pt, mp, m);
    ++g_launch_count;
    for (int j = 0; j < kNb; ++j) {
        pivot_step_kernel<<<1, 1024, 0, s>>>(pt, mp, m, j, k0, ipiv, status);
        const int rows = m - (j + 1);
        if (rows > 0 && j + 1 < kNb) panel_update_kernel<<<(rows + 255) / 256, 256, 0, s>>>(pt, mp, m, j);
        g_launch_count += 2;
    }
    panel_unpack_kernel<<<tgrid, dim3(32, 8), 0, s>>>(wp, ld, pt, mp, m, la_hi, la_lo, ldla);
    ++g_launch_count;
    const int right = ld - (k0 + kNb);
    if (right > 0) {
        laswp_kernel<<<(right + 255) / 256, 256, 0, s>>>(W + k0 + kNb, ld, right, ipiv, k0);
        ++g_launch_count;
    }
}

// The interchanges of panel k0 applied to the columns LEFT of it (the L columns factored earlier).  The elimination
// itself never needs this -- like the reference it only swaps row tails (gauss_external_input.c:139-146) -- but a
// solve with the stored factors for a NEW right-hand side does: afterwards P A = L U holds with L in final row order.
void launch_laswp_left(float* W, int ld, int k0, const int* ipiv, cudaStream_t s)
{
    if (k0 <= 0) return;
    laswp_kernel<<<(k0 + 255) / 256, 256, 0, s>>>(W, ld, k0, ipiv, k0);
    ++g_launch_count;
}

// v <- P v: every interchange of the factorisation, in pivot order, applied to a vector of np floats.  One CTA; the
// vector sits in shared memory while one thread walks the pivot list (the swaps are inherently sequential).
__global__ void __launch_bounds__(1024)
apply_pivots_vec_kernel(float* __restrict__ v, const int* __restrict__ ipiv, int np)
{
    extern __shared__ float sv[];
    for (int i = threadIdx.x; i < np; i += blockDim.x) sv[i] = v[i];
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 0; k < np; ++k) {
            const int p = ipiv[k];
            if (p != k) { const float a = sv[k]; sv[k] = sv[p]; sv[p] = a; }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < np; i += blockDim.x) v[i] = sv[i];
}
int launch_apply_pivots_vec(float* v, const int* ipiv, int np, cudaStream_t s)
{
    const size_t bytes = (size_t)np * sizeof(float);
    if (bytes > 200 * 1024) return 1;                       // the caller reports the limit
    static bool configured[16] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 16 && !configured[dev]) {
        cudaFuncSetAttribute(apply_pivots_vec_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        configured[dev] = true;
    }
    apply_pivots_vec_kernel<<<1, 1024, bytes, s>>>(v, ipiv, np);
    ++g_launch_count;
    return 0;
}

} // namespace gb200
