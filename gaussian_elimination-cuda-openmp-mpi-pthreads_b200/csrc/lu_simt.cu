// SIMT (FP32 FFMA) kernels of the blocked no-pivot LU: the FFMA GEMM used by GAUSS_CUDA_MODE_FFMA,
// the blocked back substitution and a few layout helpers (the panel kernels are in lu_panel.cu).
// sm_100a only.
//
// Reference semantics being reproduced (OpenMP_and_MPI/gauss_openmp/gauss_internal_input.c):
//   getPivot  :100-111  row i is divided by its pivot (true division), diagonal becomes 1
//   computeGauss :129-136  rows j>i:  m = a_ji ; a_j[i+1:] -= m * a_i[i+1:] ; b_j -= m * b_i
//   solveGauss :144-159  unit-diagonal back substitution
// i.e. Crout LU: A = L*U, U unit upper, L carries the pivots; the RHS is carried as an extra
// matrix column so `b_j -= m*b_i` (:135) needs no code of its own.
#include "gauss_internal.cuh"
#include <stdio.h>

namespace gb200 {

thread_local long long g_launch_count = 0;
thread_local bool t_pdl = false;

__device__ __forceinline__ float tf32_hi(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
__device__ __forceinline__ void tf32_split(float x, float& hi, float& lo) {
    hi = tf32_hi(x);
    lo = tf32_hi(x - hi);
}

// (1-3: the diagonal-block factor and the two panel solves live in lu_panel.cu)

// =====================================================================================
// 4. FP32 FFMA GEMM, NN, row-major: the pure-FFMA variant of the trailing update
//    (GAUSS_CUDA_MODE_FFMA) and of matmul_cuda.  128x128x16 CTA tile, 256 threads,
//    8x8 register tile per thread, double-buffered shared memory.
// =====================================================================================
constexpr int GBM = 128, GBN = 128, GBK = 16, GAS = GBM + 4;

__global__ void __launch_bounds__(256, 2)
sgemm_nn_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
                float* __restrict__ C, int ldc, int m, int n, int k, float alpha, float beta)
{
    __shared__ __align__(16) float As[2][GBK][GAS];   // As[k][m] (transposed on the way in)
    __shared__ __align__(16) float Bs[2][GBK][GBN];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * GBM, n0 = blockIdx.x * GBN;

    // global -> register staging coordinates
    const int a_r = tid >> 2, a_c = (tid & 3) << 2;          // rows a_r, a_r+64 ; cols a_c..a_c+3
    const int b_r = tid >> 5, b_c = (tid & 31) << 2;          // rows b_r, b_r+8  ; cols b_c..b_c+3
    const bool a_ok0 = (m0 + a_r) < m, a_ok1 = (m0 + a_r + 64) < m;
    const bool b_ok = (n0 + b_c) < n;
    const float* Ap0 = A + (size_t)(m0 + a_r) * lda + a_c;
    const float* Ap1 = A + (size_t)(m0 + a_r + 64) * lda + a_c;
    const float* Bp0 = B + (size_t)b_r * ldb + n0 + b_c;
    const float* Bp1 = B + (size_t)(b_r + 8) * ldb + n0 + b_c;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);

    float4 ra0 = a_ok0 ? *reinterpret_cast<const float4*>(Ap0) : z4;
    float4 ra1 = a_ok1 ? *reinterpret_cast<const float4*>(Ap1) : z4;
    float4 rb0 = b_ok ? *reinterpret_cast<const float4*>(Bp0) : z4;
    float4 rb1 = b_ok ? *reinterpret_cast<const float4*>(Bp1) : z4;

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    auto stage = [&](int buf) {
        As[buf][a_c][a_r] = ra0.x; As[buf][a_c + 1][a_r] = ra0.y; As[buf][a_c + 2][a_r] = ra0.z; As[buf][a_c + 3][a_r] = ra0.w;
        As[buf][a_c][a_r + 64] = ra1.x; As[buf][a_c + 1][a_r + 64] = ra1.y; As[buf][a_c + 2][a_r + 64] = ra1.z; As[buf][a_c + 3][a_r + 64] = ra1.w;
        *reinterpret_cast<float4*>(&Bs[buf][b_r][b_c]) = rb0;
        *reinterpret_cast<float4*>(&Bs[buf][b_r + 8][b_c]) = rb1;
    };
    stage(0);
    __syncthreads();

    const int nk = k / GBK;
    for (int kt = 0; kt < nk; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < nk) {
            const size_t ao = (size_t)(kt + 1) * GBK;
            const size_t bo = (size_t)(kt + 1) * GBK * ldb;
            ra0 = a_ok0 ? *reinterpret_cast<const float4*>(Ap0 + ao) : z4;
            ra1 = a_ok1 ? *reinterpret_cast<const float4*>(Ap1 + ao) : z4;
            rb0 = b_ok ? *reinterpret_cast<const float4*>(Bp0 + bo) : z4;
            rb1 = b_ok ? *reinterpret_cast<const float4*>(Bp1 + bo) : z4;
        }
#pragma unroll
        for (int kk = 0; kk < GBK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (kt + 1 < nk) {
            stage(buf ^ 1);
            __syncthreads();
        }
    }

    // epilogue: C = beta*C + alpha*acc, float4, row/column guarded
#pragma unroll
    for (int ih = 0; ih < 2; ++ih)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = m0 + ih * 64 + ty * 4 + i;
            if (r >= m) continue;
#pragma unroll
            for (int jh = 0; jh < 2; ++jh) {
                const int cc = n0 + jh * 64 + tx * 4;
                if (cc >= n) continue;
                float* cp = C + (size_t)r * ldc + cc;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (beta != 0.f) v = *reinterpret_cast<const float4*>(cp);
                const float* ac = acc[ih * 4 + i] + jh * 4;
                v.x = fmaf(alpha, ac[0], v.x); v.y = fmaf(alpha, ac[1], v.y);
                v.z = fmaf(alpha, ac[2], v.z); v.w = fmaf(alpha, ac[3], v.w);
                *reinterpret_cast<float4*>(cp) = v;
            }
        }
}

void launch_sgemm_nn(const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                     int m, int n, int k, float alpha, float beta, cudaStream_t s)
{
    if (m <= 0 || n <= 0 || k <= 0) return;
    dim3 grid((n + GBN - 1) / GBN, (m + GBM - 1) / GBM);
    sgemm_nn_kernel<<<grid, 256, 0, s>>>(A, lda, B, ldb, C, ldc, m, n, k, alpha, beta);
    ++g_launch_count;
}

// =====================================================================================
// 5. Back substitution (solveGauss :144-159) in ONE launch.  CTAs take 128-row blocks bottom-up through
//     an atomic ticket (so a CTA only ever waits for CTAs that already run), stream their rows of U
//     with float4 loads while folding in x_J as the blocks below publish it (per-block flags,
//     release/acquire), then solve their own diagonal block with the inverses of its four 32x32
//     sub-blocks and publish x_r.  U is read exactly once (2*N^2 bytes in total); the dependent chain
//     per block is: flag -> one 128x128 block product -> 4 small matvecs -> flag.
// =====================================================================================
constexpr int TFL = kNb + 4;                 // smem row stride of the diagonal block
struct TrsvSmem {
    float U[kNb][TFL];                       // this CTA's diagonal block
    float invT[4][32][33];                   // invT[b][k][i] = (U_bb^-1)[i][k]
    float v[kNb];                            // running right-hand side of the block, finally x
    float red[kNb];
    int ticket;
};

__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// LOWER = false: x = U^-1 y, U unit upper (solveGauss).  LOWER = true: y = L^-1 r, L lower with the pivots on its
// diagonal (the forward half of a solve with the stored factors, used by the iterative-refinement entry point).
// rhs != nullptr: the right-hand side is that vector instead of column `ycol` of W.
template <bool LOWER>
__global__ void __launch_bounds__(256, 1)
trsv_fused_kernel(const float* __restrict__ W, int ld, int nrows, int ycol, const float* __restrict__ rhs,
                  float* __restrict__ x, int* __restrict__ ctrl)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TrsvSmem& sm = *reinterpret_cast<TrsvSmem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nblk = nrows / kNb;
    if (tid == 0) sm.ticket = atomicAdd(&ctrl[0], 1);
    __syncthreads();
    // my row block; every block it depends on belongs to an earlier ticket (so it is running or done)
    const int r = LOWER ? sm.ticket : nblk - 1 - sm.ticket;
    int* flags = ctrl + 1;
    const float* Wr = W + (size_t)r * kNb * ld;

    // own diagonal block -> smem, its RHS entries -> v
    for (int idx = tid; idx < kNb * kNb / 4; idx += 256) {
        const int rr = idx >> 5, c4 = (idx & 31) << 2;
        *reinterpret_cast<float4*>(&sm.U[rr][c4]) = __ldg(reinterpret_cast<const float4*>(Wr + (size_t)rr * ld + r * kNb + c4));
    }
    if (tid < kNb) sm.v[tid] = rhs ? rhs[r * kNb + tid] : Wr[(size_t)tid * ld + ycol];

    // rows 16*warp .. 16*warp+15 of the block; per-lane partial dot products over the lane's 4 columns
    float part[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) part[i] = 0.f;
    const float* rowbase = Wr + (size_t)(16 * warp) * ld + 4 * lane;
    float4 cur[16];
    const int Jfirst = LOWER ? 0 : nblk - 1, Jstep = LOWER ? 1 : -1;
    const int nJ = LOWER ? r : nblk - 1 - r;                 // off-diagonal blocks of this block row
    if (nJ > 0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) cur[i] = __ldcs(reinterpret_cast<const float4*>(rowbase + (size_t)i * ld + Jfirst * kNb));
    }
    __syncthreads();
    if (warp < 4) {
        // inverse of the 32x32 diagonal sub-block `warp`: lane c solves for column c of the inverse in registers
        const int o = 32 * warp;
        float xi[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) xi[i] = (i == lane) ? 1.f : 0.f;
        if (LOWER) {
            const float myrinv = 1.0f / sm.U[o + lane][o + lane];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                xi[j] *= __shfl_sync(0xffffffffu, myrinv, j);
#pragma unroll
                for (int i = j + 1; i < 32; ++i) xi[i] = fmaf(-sm.U[o + i][o + j], xi[j], xi[i]);
            }
        } else {
#pragma unroll
            for (int j = 31; j > 0; --j) {
#pragma unroll
                for (int i = 0; i < j; ++i) xi[i] = fmaf(-sm.U[o + i][o + j], xi[j], xi[i]);
            }
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) sm.invT[warp][lane][i] = xi[i];
    }

    for (int jj = 0, J = Jfirst; jj < nJ; ++jj, J += Jstep) {
        float4 nxt[16];
        if (jj + 1 < nJ) {                    // prefetch the next block before waiting for x_J
#pragma unroll
            for (int i = 0; i < 16; ++i) nxt[i] = __ldcs(reinterpret_cast<const float4*>(rowbase + (size_t)i * ld + (J + Jstep) * kNb));
        }
        if (lane == 0) {                      // bounded wait: a protocol bug must trap, never hang the GPU
            const long long t0 = clock64();
            while (ld_acquire(flags + J) == 0) {
                __nanosleep(64);
                if (clock64() - t0 > 4000000000LL) { printf("trsv_fused: wait for block %d timed out (block %d)\n", J, r); __trap(); }
            }
        }
        __syncwarp();
        const float4 xv = __ldcg(reinterpret_cast<const float4*>(x + J * kNb + 4 * lane));
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            part[i] = fmaf(cur[i].x, xv.x, part[i]); part[i] = fmaf(cur[i].y, xv.y, part[i]);
            part[i] = fmaf(cur[i].z, xv.z, part[i]); part[i] = fmaf(cur[i].w, xv.w, part[i]);
        }
        if (jj + 1 < nJ) {
#pragma unroll
            for (int i = 0; i < 16; ++i) cur[i] = nxt[i];
        }
    }
    // fold the per-lane partial sums: v[row] -= sum over lanes
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        float s_ = part[i];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) s_ += __shfl_xor_sync(0xffffffffu, s_, off);
        if (lane == 0) sm.red[16 * warp + i] = s_;
    }
    __syncthreads();
    if (tid < kNb) sm.v[tid] -= sm.red[tid];
    __syncthreads();
    // diagonal block: x_b = inv(D_bb) * v_b, then the other sub-blocks' v -= (off-diagonal sub-block) * x_b;
    // b runs 3..0 for the upper solve and 0..3 for the lower one
    for (int bi = 0; bi < 4; ++bi) {
        const int b = LOWER ? bi : 3 - bi;
        if (warp == b) {
            float acc = 0.f;
#pragma unroll 8
            for (int k = 0; k < 32; ++k) acc = fmaf(sm.invT[b][k][lane], sm.v[32 * b + k], acc);
            __syncwarp();
            sm.v[32 * b + lane] = acc;
        }
        __syncthreads();
        if (warp < 4 && (LOWER ? warp > b : warp < b)) {
            float acc = sm.v[tid];
#pragma unroll 8
            for (int k = 0; k < 32; ++k) acc = fmaf(-sm.U[tid][32 * b + k], sm.v[32 * b + k], acc);
            sm.v[tid] = acc;
        }
        __syncthreads();
    }
    if (tid < kNb) x[r * kNb + tid] = sm.v[tid];
    __threadfence();
    __syncthreads();
    if (tid == 0) st_release(flags + r, 1);
}

void launch_trsv_fused(const float* W, int ld, int nrows, int ycol, float* x, int* ctrl, cudaStream_t s)
{
    const int nblk = nrows / kNb;
    cudaMemsetAsync(ctrl, 0, (size_t)(nblk + 1) * sizeof(int), s);
    trsv_fused_kernel<false><<<nblk, 256, sizeof(TrsvSmem), s>>>(W, ld, nrows, ycol, nullptr, x, ctrl);
    ++g_launch_count;
}

// Solve with the stored factors for a new right-hand side (single GPU):  y = L^-1 rhs  (lower = true) or
// x = U^-1 rhs (lower = false); rhs and out are device vectors of nrows floats, W is not modified.
void launch_trsv_factor_solve(const float* W, int ld, int nrows, const float* rhs, float* out, bool lower, int* ctrl, cudaStream_t s)
{
    const int nblk = nrows / kNb;
    cudaMemsetAsync(ctrl, 0, (size_t)(nblk + 1) * sizeof(int), s);
    if (lower) trsv_fused_kernel<true><<<nblk, 256, sizeof(TrsvSmem), s>>>(W, ld, nrows, 0, rhs, out, ctrl);
    else       trsv_fused_kernel<false><<<nblk, 256, sizeof(TrsvSmem), s>>>(W, ld, nrows, 0, rhs, out, ctrl);
    ++g_launch_count;
}

// ---- FP64 helpers of the iterative-refinement entry point --------------------------------------
// W[i][j] = (float)A[i][j] for i, j < n (A row-major n x n doubles), RHS column `ycol` = (float)b[i]
__global__ void f64_to_w_kernel(const double* __restrict__ A, const double* __restrict__ b, int n, float* __restrict__ w, int ld, int ycol)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j < n) w[(size_t)i * ld + j] = (float)A[(size_t)i * n + j];
    if (j == 0) w[(size_t)i * ld + ycol] = (float)b[i];
}
// r = b - A x in double (a warp per row), rf[i] = (float)r[i]; rows n..np-1 of rf are zeroed by the caller
__global__ void __launch_bounds__(256)
residual_f64_kernel(const double* __restrict__ A, const double* __restrict__ x, const double* __restrict__ b,
                    double* __restrict__ r, float* __restrict__ rf, int n)
{
    const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= n) return;
    const double* a = A + (size_t)row * n;
    double acc = 0.0;
    for (int j = lane; j < n; j += 32) acc = fma(a[j], x[j], acc);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if (lane == 0) { const double rr = b[row] - acc; r[row] = rr; rf[row] = (float)rr; }
}
// x += d (d float), and the two norms the stopping test needs: out[0] = max|d|, out[1] = max|x| (one block)
__global__ void __launch_bounds__(1024)
refine_update_kernel(double* __restrict__ x, const float* __restrict__ d, int n, double* __restrict__ out)
{
    __shared__ double sd[32], sx[32];
    double md = 0.0, mx = 0.0;
    for (int i = threadIdx.x; i < n; i += 1024) {
        const double di = (double)d[i];
        const double xi = x[i] + di;
        x[i] = xi;
        md = fmax(md, fabs(di)); mx = fmax(mx, fabs(xi));
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        md = fmax(md, __shfl_xor_sync(0xffffffffu, md, off));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    }
    if ((threadIdx.x & 31) == 0) { sd[threadIdx.x >> 5] = md; sx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x < 32) {
        md = sd[threadIdx.x]; mx = sx[threadIdx.x];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            md = fmax(md, __shfl_xor_sync(0xffffffffu, md, off));
            mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, off));
        }
        if (threadIdx.x == 0) { out[0] = md; out[1] = mx; }
    }
}
void launch_f64_to_w(const double* A, const double* b, int n, float* w, int ld, int ycol, cudaStream_t s)
{
    dim3 grid((n + 255) / 256, n);
    f64_to_w_kernel<<<grid, 256, 0, s>>>(A, b, n, w, ld, ycol);
    ++g_launch_count;
}
void launch_residual_f64(const double* A, const double* x, const double* b, double* r, float* rf, int n, cudaStream_t s)
{
    residual_f64_kernel<<<(n + 7) / 8, 256, 0, s>>>(A, x, b, r, rf, n);
    ++g_launch_count;
}
void launch_refine_update(double* x, const float* d, int n, double* out, cudaStream_t s)
{
    refine_update_kernel<<<1, 1024, 0, s>>>(x, d, n, out);
    ++g_launch_count;
}

// y[r] -= U[r, 0..kw) . x[0..kw)  for nrows rows (kw a multiple of 128): the wide form of trsv_gemv, one
// launch per outer block in the multi-GPU back substitution.  A warp takes 4 rows, float4 per lane.
__global__ void __launch_bounds__(256)
trsv_gemv_wide_kernel(float* __restrict__ u, int ld, int ycol, const float* __restrict__ x, int nrows, int kw)
{
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int r0 = (blockIdx.x * 8 + wrp) * 4;
    float s_[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < kw; c += kNb) {
        const float4 xv = __ldg(reinterpret_cast<const float4*>(x + c + 4 * lane));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (r0 + i < nrows) {
                const float4 v = __ldcs(reinterpret_cast<const float4*>(u + (size_t)(r0 + i) * ld + c + 4 * lane));
                s_[i] = fmaf(v.x, xv.x, s_[i]); s_[i] = fmaf(v.y, xv.y, s_[i]);
                s_[i] = fmaf(v.z, xv.z, s_[i]); s_[i] = fmaf(v.w, xv.w, s_[i]);
            }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1)
#pragma unroll
        for (int i = 0; i < 4; ++i) s_[i] += __shfl_xor_sync(0xffffffffu, s_[i], off);
    if (lane < 4 && r0 + lane < nrows) {
        const float mine = lane == 0 ? s_[0] : lane == 1 ? s_[1] : lane == 2 ? s_[2] : s_[3];
        u[(size_t)(r0 + lane) * ld + ycol] -= mine;
    }
}

void launch_trsv_gemv_wide(float* u, int ld, int ycol, const float* x, int nrows, int kw, cudaStream_t s)
{
    if (nrows <= 0 || kw <= 0) return;
    trsv_gemv_wide_kernel<<<(nrows + 31) / 32, 256, 0, s>>>(u, ld, ycol, x, nrows, kw);
    ++g_launch_count;
}

// =====================================================================================
// 6. Layout helpers
// =====================================================================================
__global__ void set_column_kernel(float* w, int ld, int col, const float* src, int nrows)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < nrows) w[(size_t)r * ld + col] = src[r];
}
__global__ void get_column_kernel(const float* w, int ld, int col, float* dst, int nrows)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < nrows) dst[r] = w[(size_t)r * ld + col];
}
// Rows/columns beyond n are identity: [[A 0][0 I]], rhs 0 -- the padded unknowns come out 0 and
// never touch the real ones.  One thread per (row, column) over columns [col0, ld) of `nrows` local rows;
// real matrix entries and real RHS entries (column ld - kPadCols) are left alone.
__global__ void pad_identity_kernel(float* w, int ld, int row0_local, int row0_global, int nrows, int n, int col0)
{
    const int c = col0 + blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y;
    if (c >= ld || r >= nrows) return;
    const int g = row0_global + r;
    if (g < n && (c < n || c == ld - kPadCols)) return;
    w[(size_t)(row0_local + r) * ld + c] = (g >= n && c == g) ? 1.f : 0.f;
}

// initMatrix on the device (gauss_internal_input.c:50-60): matrix[i][j] = 2 (min(i, j) + 1), B[i] = i, for `nrows`
// local rows starting at global row g0; the RHS lives in column ycol.  Rows / columns beyond n are left to pad_identity.
__global__ void init_fixture_kernel(float* __restrict__ w, int ld, int row0_local, int g0, int nrows, int n, int ycol)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y;
    const int i = g0 + r;
    if (r >= nrows || i >= n) return;
    float* row = w + (size_t)(row0_local + r) * ld;
    if (j < n) row[j] = 2.0f * (float)((j < i ? j : i) + 1);
    if (j == 0) row[ycol] = (float)i;
}
void launch_init_fixture(float* w, int ld, int row0_local, int g0, int nrows, int n, int ycol, cudaStream_t s)
{
    if (nrows <= 0) return;
    dim3 grid((n + 255) / 256, nrows);
    init_fixture_kernel<<<grid, 256, 0, s>>>(w, ld, row0_local, g0, nrows, n, ycol);
    ++g_launch_count;
}

void launch_set_column(float* w, int ld, int col, const float* src, int nrows, cudaStream_t s)
{
    if (nrows <= 0) return;
    set_column_kernel<<<(nrows + 255) / 256, 256, 0, s>>>(w, ld, col, src, nrows);
    ++g_launch_count;
}
void launch_get_column(const float* w, int ld, int col, float* dst, int nrows, cudaStream_t s)
{
    if (nrows <= 0) return;
    get_column_kernel<<<(nrows + 255) / 256, 256, 0, s>>>(w, ld, col, dst, nrows);
    ++g_launch_count;
}
void launch_pad_identity(float* w, int ld, int row0_local, int row0_global, int nrows, int n, int col0, cudaStream_t s)
{
    if (nrows <= 0 || col0 >= ld) return;
    dim3 grid((ld - col0 + 255) / 256, nrows);
    pad_identity_kernel<<<grid, 256, 0, s>>>(w, ld, row0_local, row0_global, nrows, n, col0);
    ++g_launch_count;
}

// ---- matmul operand preparation (tensor-core path) ----
// hi/lo[r][c] = split(src[r][c]) for r < rows, c < ldd; columns >= cols are zero (K padding).
__global__ void split_rowmajor_kernel(const float* __restrict__ src, int lds, float* __restrict__ hi, float* __restrict__ lo,
                                      int ldd, int rows, int cols)
{
    const int c4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int r = blockIdx.y;
    pdl_release();
    pdl_wait();
    if (c4 >= ldd || r >= rows) return;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 4; ++i) if (c4 + i < cols) v[i] = src[(size_t)r * lds + c4 + i];
    float4 h, l;
    tf32_split(v[0], h.x, l.x); tf32_split(v[1], h.y, l.y); tf32_split(v[2], h.z, l.z); tf32_split(v[3], h.w, l.w);
    *reinterpret_cast<float4*>(hi + (size_t)r * ldd + c4) = h;
    *reinterpret_cast<float4*>(lo + (size_t)r * ldd + c4) = l;
}
// hi/lo[c*ldd + r] = split(src[r][c]) for c < cols, r < rext; rows >= `rows` of the source read as zero.
// 32x32 tiles through shared memory, both sides coalesced.
__global__ void split_transpose_kernel(const float* __restrict__ src, int lds, float* __restrict__ hi, float* __restrict__ lo,
                                       int ldd, int rows, int cols, int rext)
{
    __shared__ float t[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;          // 32 x 8
    pdl_release();
    pdl_wait();
    for (int i = ty; i < 32; i += 8) {
        const int r = r0 + i, c = c0 + tx;
        t[i][tx] = (r < rows && c < cols) ? src[(size_t)r * lds + c] : 0.f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, r = r0 + tx;                 // output row = source column
        if (c >= cols || r >= rext) continue;
        float h, l;
        tf32_split(t[tx][i], h, l);
        hi[(size_t)c * ldd + r] = h;
        lo[(size_t)c * ldd + r] = l;
    }
}
void launch_split_rowmajor(const float* src, int lds, float* hi, float* lo, int ldd, int rows, int cols, cudaStream_t s)
{
    dim3 grid((ldd / 4 + 127) / 128, rows);
    launch_k(split_rowmajor_kernel, grid, dim3(128), 0, s, src, lds, hi, lo, ldd, rows, cols);
    ++g_launch_count;
}
void launch_split_transpose(const float* src, int lds, float* hi, float* lo, int ldd, int rows, int cols, int rext, cudaStream_t s)
{
    if (cols <= 0 || rext <= 0) return;
    dim3 grid((cols + 31) / 32, (rext + 31) / 32);
    launch_k(split_transpose_kernel, grid, dim3(32, 8), 0, s, src, lds, hi, lo, ldd, rows, cols, rext);
    ++g_launch_count;
}

// Dynamic shared-memory opt-ins are per device: call once per context with that device current.
void simt_setup_device()
{
    panel_setup_device();
    cudaFuncSetAttribute(trsv_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TrsvSmem));
    cudaFuncSetAttribute(trsv_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TrsvSmem));
}

} // namespace gb200
