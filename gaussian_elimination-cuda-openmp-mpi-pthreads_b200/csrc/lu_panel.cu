// Panel kernels of the blocked no-pivot LU (the latency-critical part of computeGauss,
// gauss_internal_input.c:118-138): the 128x128 diagonal-block factor and the two triangular
// panel solves.  The solves work on 32-wide sub-blocks; their block products run as FP32 FFMAs
// (sections 2-3, FFMA mode) or as warp-level TF32 MMAs (section 4, 3xTF32 mode).
//
//   diag_factor  A11 = L11*U11 in registers, two pivots per barrier (Crout: U unit upper, L carries the
//                pivots; getPivot :100-111 scales row i by 1/pivot, computeGauss :129-136 eliminates)
//   urow         U12 = L11^-1 * A12      (the part of rows i right of the diagonal block, :132-134)
//   lcol         L21 = A21 * U11^-1      (the multipliers a_ji of the rows below, :130)
// urow/lcol invert the four 32x32 diagonal sub-blocks once per CTA (one warp each) and then run
// block forward substitution:  X_b = inv(D_b) * (A_b - sum_{c<b} L_bc X_c)  as dense products.
#include "gauss_internal.cuh"
#include <type_traits>

namespace gb200 {

__device__ __forceinline__ float tf32_round(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
__device__ __forceinline__ void tf32_split2(float x, float& hi, float& lo) {
    hi = tf32_round(x);
    lo = tf32_round(x - hi);
}
// 16-byte asynchronous global -> shared copy (LDGSTS): all copies of a stage are in flight together, so
// the global-memory latency is paid once per stage instead of once per loop iteration
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// =====================================================================================
// 1. Diagonal block factor: 128x128, one CTA of 256 threads, the block lives in registers
//    (16x16 thread grid, cyclic 8x8 elements per thread).  Latency is everything here, so the
//    per-pivot chain is kept to  STS -> barrier -> LDS -> rcp -> FMUL -> FFMA :
//      * row k and column k are published RAW (no normalisation on the chain); every thread forms
//        the reciprocal of the pivot itself and scales its multipliers (a_ik / a_kk), i.e. the step is
//        done in Doolittle form and U is scaled to the reference's unit-diagonal form (getPivot
//        :100-111) once at the end, with the same reciprocals;
//      * the update is branch-free (masked operands instead of predicated code), the loops over
//        the 16 pivots of a phase are rolled so the code stays hot in the instruction cache.
// =====================================================================================
__device__ __forceinline__ float fast_rcp(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return fmaf(r, fmaf(-x, r, 1.0f), r);                 // one Newton step: full FP32 accuracy, exact for powers of two
}
// A pivot the elimination cannot divide by: zero, NaN, infinite, subnormal (rcp.approx.ftz flushes it to zero and
// returns inf) or so large that its reciprocal is flushed to zero.  `r` is fast_rcp(p).
__device__ __forceinline__ bool bad_pivot(float p, float r)
{
    return !(fabsf(p) >= 1.17549435e-38f) || !(fabsf(r) >= 1.17549435e-38f) || isinf(p) || isinf(r);
}

struct DiagSmem2 {
    float rowbuf[2][2][kNb];
    float colbuf[2][2][kNb];
    float rinv[kNb];
    int first_bad;
};

// Two pivots per barrier.  Rows k, k+1 and columns k, k+1 are published raw; every thread then redoes the tiny
// part of step k that step k+1 depends on (the 2x2 pivot block, its own entries of row k+1 and column k+1) and
// applies both rank-1 updates back to back.  Each element sees exactly the operations of two single steps in the
// same order; the barrier / publish / reload latency is paid once per two pivots.
template <int PH>
__device__ __forceinline__ void diag_phase2(float (&a)[8][8], DiagSmem2& sm, const int tx, const int ty)
{
#pragma unroll 1
    for (int kk = 0; kk < 16; kk += 2) {
        const int k = PH * 16 + kk;
        const int par = (kk >> 1) & 1;
        float* rb0 = sm.rowbuf[par][0];
        float* rb1 = sm.rowbuf[par][1];
        float* cb0 = sm.colbuf[par][0];
        float* cb1 = sm.colbuf[par][1];
        if (ty == kk || ty == kk + 1) {                    // holders of rows k, k+1
            float* rb = (ty == kk) ? rb0 : rb1;
#pragma unroll
            for (int q = PH; q < 8; ++q) rb[tx + 16 * q] = a[PH][q];
        }
        if (tx == kk || tx == kk + 1) {                    // holders of columns k, k+1
            float* cb = (tx == kk) ? cb0 : cb1;
#pragma unroll
            for (int p = PH; p < 8; ++p) cb[ty + 16 * p] = a[p][PH];
        }
        __syncthreads();
        const float p0 = rb0[k], b01 = rb0[k + 1], c10 = rb1[k], d11 = rb1[k + 1];
        const float r0 = fast_rcp(p0);
        const float m10 = c10 * r0;                        // multiplier of row k+1 in step k
        const float p1 = fmaf(-m10, b01, d11);             // pivot of step k+1
        const float r1 = fast_rcp(p1);
        if (tx == kk && ty == kk) {
            sm.rinv[k] = r0;
            sm.rinv[k + 1] = r1;
            if (bad_pivot(p0, r0)) atomicMin(&sm.first_bad, k);
            else if (bad_pivot(p1, r1)) atomicMin(&sm.first_bad, k + 1);
        }
        // only the first row/column group of the phase can still contain indices <= k+1
        float u0[8], u1[8], l0[8], l1[8];
        {
            const float x0 = rb0[tx + 16 * PH], x1 = rb1[tx + 16 * PH];
            u0[PH] = (tx > kk) ? x0 : 0.f;
            u1[PH] = (tx > kk + 1) ? fmaf(-m10, x0, x1) : 0.f;
            const float y0 = cb0[ty + 16 * PH], y1 = cb1[ty + 16 * PH];
            l0[PH] = (ty > kk) ? y0 * r0 : 0.f;
            l1[PH] = (ty > kk + 1) ? fmaf(-l0[PH], b01, y1) * r1 : 0.f;
        }
#pragma unroll
        for (int q = PH + 1; q < 8; ++q) {
            u0[q] = rb0[tx + 16 * q];
            u1[q] = fmaf(-m10, u0[q], rb1[tx + 16 * q]);
        }
#pragma unroll
        for (int p = PH + 1; p < 8; ++p) {
            l0[p] = cb0[ty + 16 * p] * r0;
            l1[p] = fmaf(-l0[p], b01, cb1[ty + 16 * p]) * r1;
        }
#pragma unroll
        for (int p = PH; p < 8; ++p)
#pragma unroll
            for (int q = PH; q < 8; ++q) {
                a[p][q] = fmaf(-l0[p], u0[q], a[p][q]);     // computeGauss :132-134, step k
                a[p][q] = fmaf(-l1[p], u1[q], a[p][q]);     //                        step k+1
            }
    }
}

// The kernel: two pivots per barrier.  Measured on B200: 23.9 us against 27.3 us for one pivot per barrier (a variant kept
// through round 1 for A/B timing, bit-identical output).  (Four pivots per barrier: 24.1 us, eight:
// 52 us -- the redundant work grows faster than the barrier count shrinks.  A variant blocked 32 wide with the
// 32 sequential pivots inside one warp, rows exchanged by shuffle, took 52 us: a lone warp issues its
// dependent shuffle/FMA stream at ~3 cycles per instruction.)
__global__ void __launch_bounds__(256, 1)
diag_factor_pair_kernel(float* __restrict__ blk, int ld, int grow0, int* __restrict__ status)
{
    __shared__ DiagSmem2 sm;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    pdl_release();
    if (tid == 0) sm.first_bad = kNb;
    pdl_wait();
    float a[8][8];
#pragma unroll
    for (int p = 0; p < 8; ++p)
#pragma unroll
        for (int q = 0; q < 8; ++q) a[p][q] = blk[(size_t)(ty + 16 * p) * ld + tx + 16 * q];
    diag_phase2<0>(a, sm, tx, ty);
    diag_phase2<1>(a, sm, tx, ty);
    diag_phase2<2>(a, sm, tx, ty);
    diag_phase2<3>(a, sm, tx, ty);
    diag_phase2<4>(a, sm, tx, ty);
    diag_phase2<5>(a, sm, tx, ty);
    diag_phase2<6>(a, sm, tx, ty);
    diag_phase2<7>(a, sm, tx, ty);
    __syncthreads();
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        const int i = ty + 16 * p;
        const float ri = sm.rinv[i];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int j = tx + 16 * q;
            const float v = (j > i) ? a[p][q] * ri : a[p][q];
            blk[(size_t)i * ld + j] = v;
        }
    }
    if (tid == 0 && sm.first_bad < kNb) atomicCAS(status, 0, grow0 + sm.first_bad + 1);
}

void launch_diag_factor(float* blk, int ld, int global_row0, int* status, cudaStream_t s)
{
    launch_k(diag_factor_pair_kernel, dim3(1), dim3(256), 0, s, blk, ld, global_row0, status);
    ++g_launch_count;
}

// =====================================================================================
// 2. U-row panel:  X = L11^-1 * A12, 64 matrix columns per CTA, 256 threads.
//    Code is kept compact on purpose (rolled loops, one shared multiply-accumulate body): these
//    CTAs run each instruction only a few times, so straight-line unrolled code is bound by cold
//    instruction fetch rather than by arithmetic.
// =====================================================================================
constexpr int UW = 64;
struct UrowSmem {
    float L[kNb][kNb];            // row-major L11 (i >= j part used)
    float invLd[4][32][32];       // inverses of the four 32x32 diagonal sub-blocks of L11, row-major
    float strip[kNb][UW];
};

// acc[i][c] += sum_{k<32} A[i][k] * B[k][lane + 32c]   for 4 rows of A (row stride lda) and a 32 x 64 block B
__device__ __forceinline__ void mac_4x2(float (&acc)[4][2], const float* __restrict__ A, const int lda,
                                        const float* __restrict__ B, const int lane)
{
#pragma unroll 2
    for (int k4 = 0; k4 < 32; k4 += 4) {
        float4 a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(A + i * lda + k4);
        float bv[4][2];
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) { bv[kk][0] = B[(k4 + kk) * UW + lane]; bv[kk][1] = B[(k4 + kk) * UW + lane + 32]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            acc[i][0] = fmaf(a[i].x, bv[0][0], acc[i][0]); acc[i][1] = fmaf(a[i].x, bv[0][1], acc[i][1]);
            acc[i][0] = fmaf(a[i].y, bv[1][0], acc[i][0]); acc[i][1] = fmaf(a[i].y, bv[1][1], acc[i][1]);
            acc[i][0] = fmaf(a[i].z, bv[2][0], acc[i][0]); acc[i][1] = fmaf(a[i].z, bv[2][1], acc[i][1]);
            acc[i][0] = fmaf(a[i].w, bv[3][0], acc[i][0]); acc[i][1] = fmaf(a[i].w, bv[3][1], acc[i][1]);
        }
    }
}

__global__ void __launch_bounds__(256, 2)
urow_kernel(const float* __restrict__ l11, int ldl, float* __restrict__ a12, int ld, int ncols,
            float* __restrict__ ub_hi, float* __restrict__ ub_lo, int ldub, int k_ub)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    UrowSmem& sm = *reinterpret_cast<UrowSmem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c0 = blockIdx.x * UW;

    // stage 0: the four 32x32 diagonal sub-blocks of L11 (needed first, for the inverses)
    for (int idx = tid; idx < 4 * 32 * 8; idx += 256) {
        const int bb = idx >> 8, r = 32 * bb + ((idx >> 3) & 31), c4 = 32 * bb + ((idx & 7) << 2);
        cp_async16(&sm.L[r][c4], l11 + (size_t)r * ldl + c4);
    }
    cp_async_commit();
    // stage 1: the rest of L11 and the 128 x 64 strip (columns beyond ncols are the zero padding of W)
    for (int idx = tid; idx < kNb * kNb / 4; idx += 256) {
        const int r = idx >> 5, c4 = (idx & 31) << 2;
        if ((r >> 5) != (c4 >> 5)) cp_async16(&sm.L[r][c4], l11 + (size_t)r * ldl + c4);
    }
    {
        const int c4 = (tid & 15) << 2;
        const bool ok = c0 + c4 < ncols;
        for (int r = tid >> 4; r < kNb; r += 16) {
            if (ok) cp_async16(&sm.strip[r][c4], a12 + (size_t)r * ld + c0 + c4);
            else *reinterpret_cast<float4*>(&sm.strip[r][c4]) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    if (warp < 4) {
        // inverse of the 32x32 lower-triangular diagonal sub-block `warp`: lane c solves L x = e_c by
        // forward substitution, the column held in registers (the L entries are warp-wide broadcasts)
        const int o = 32 * warp;
        const float myrinv = 1.0f / sm.L[o + lane][o + lane];
        float x[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = (i == lane) ? 1.f : 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            x[j] *= __shfl_sync(0xffffffffu, myrinv, j);
#pragma unroll
            for (int i = j + 1; i < 32; ++i) x[i] = fmaf(-sm.L[o + i][o + j], x[j], x[i]);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) sm.invLd[warp][i][lane] = x[i];
    }
    cp_async_wait<0>();
    __syncthreads();

#pragma unroll 1
    for (int b = 0; b < 4; ++b) {
        const float* Bb = &sm.strip[32 * b][0];
        // X_b = inv(L_bb) * strip_b : warp w -> rows 4w..4w+3 of the sub-block, lane -> columns lane, lane+32
        float acc[4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i) { acc[i][0] = 0.f; acc[i][1] = 0.f; }
        mac_4x2(acc, &sm.invLd[b][4 * warp][0], 32, Bb, lane);
        __syncthreads();                                   // every read of strip_b is done
#pragma unroll
        for (int i = 0; i < 4; ++i) { sm.strip[32 * b + 4 * warp + i][lane] = acc[i][0]; sm.strip[32 * b + 4 * warp + i][lane + 32] = acc[i][1]; }
        __syncthreads();
        // rows below: strip[r] -= L[r, 32b..32b+31] * X_b, 32 rows per pass (4 per warp)
#pragma unroll 1
        for (int r0 = 32 * (b + 1) + 4 * warp; r0 < kNb; r0 += 32) {
#pragma unroll
            for (int i = 0; i < 4; ++i) { acc[i][0] = 0.f; acc[i][1] = 0.f; }
            mac_4x2(acc, &sm.L[r0][32 * b], kNb, Bb, lane);
#pragma unroll
            for (int i = 0; i < 4; ++i) { sm.strip[r0 + i][lane] -= acc[i][0]; sm.strip[r0 + i][lane + 32] -= acc[i][1]; }
        }
        __syncthreads();
    }

    // write back: U in place, and (3xTF32 mode) its hi/lo split transposed into the K-major B operand
    {
        const int c4 = (tid & 15) << 2;
        if (c0 + c4 < ncols)
            for (int r = tid >> 4; r < kNb; r += 16)
                *reinterpret_cast<float4*>(a12 + (size_t)r * ld + c0 + c4) = *reinterpret_cast<const float4*>(&sm.strip[r][c4]);
    }
    if (ub_hi != nullptr) {
        const int c = tid & 63, part = tid >> 6;           // 4 threads per column, 32 rows each
        if (c0 + c < ncols) {
            float* dh = ub_hi + (size_t)(c0 + c) * ldub + k_ub + 32 * part;
            float* dl = ub_lo + (size_t)(c0 + c) * ldub + k_ub + 32 * part;
#pragma unroll 2
            for (int r = 0; r < 32; r += 4) {
                float4 h, l;
                tf32_split2(sm.strip[32 * part + r][c], h.x, l.x);
                tf32_split2(sm.strip[32 * part + r + 1][c], h.y, l.y);
                tf32_split2(sm.strip[32 * part + r + 2][c], h.z, l.z);
                tf32_split2(sm.strip[32 * part + r + 3][c], h.w, l.w);
                *reinterpret_cast<float4*>(dh + r) = h;
                *reinterpret_cast<float4*>(dl + r) = l;
            }
        }
    }
}

// 3xTF32 mode runs the tensor-core variants (section 4) unless GAUSS_CUDA_PANEL_MMA=0; the FFMA mode keeps the
// pure-FP32 kernels above
static int g_panel_mma = -1;
static bool panel_mma_enabled()
{
    if (g_panel_mma < 0) { const char* e = getenv("GAUSS_CUDA_PANEL_MMA"); g_panel_mma = (e && atoi(e) == 0) ? 0 : 1; }
    return g_panel_mma != 0;
}
static int g_panel_fuse = -1;
static bool panel_fuse_enabled()           // GAUSS_CUDA_PANEL_FUSE=0: U-row and L-col solves of an inner block as two launches
{
    if (g_panel_fuse < 0) { const char* e = getenv("GAUSS_CUDA_PANEL_FUSE"); g_panel_fuse = (e && atoi(e) == 0) ? 0 : 1; }
    return g_panel_fuse != 0;
}
static void launch_urow_mma(const float* l11, int ldl, float* a12, int ld, int ncols, float* ub_hi, float* ub_lo, int ldub, int k_ub, cudaStream_t s);
static void launch_lcol_mma(const float* u11, int ldu, float* a21, int ld, int nrows, float* la_hi, float* la_lo, int ldla, int k_la, cudaStream_t s);

void launch_urow(const float* l11, int ldl, float* a12, int ld, int ncols,
                 float* ub_hi, float* ub_lo, int ldub, int k_ub, cudaStream_t s)
{
    if (ncols <= 0) return;
    if (ub_hi != nullptr && panel_mma_enabled()) launch_urow_mma(l11, ldl, a12, ld, ncols, ub_hi, ub_lo, ldub, k_ub, s);
    else urow_kernel<<<(ncols + UW - 1) / UW, 256, sizeof(UrowSmem), s>>>(l11, ldl, a12, ld, ncols, ub_hi, ub_lo, ldub, k_ub);
    ++g_launch_count;
}

// =====================================================================================
// 3. L-col panel:  X = A21 * U11^-1 (U11 unit upper), 64 matrix rows per CTA, 256 threads
// =====================================================================================
constexpr int LH = 64;
struct LcolSmem {
    float U[kNb][kNb];            // row-major U11 (j > i part used; unit diagonal implied)
    float invUd[4][32][32];       // inverses of the four 32x32 unit-upper diagonal sub-blocks
    float tile[LH][kNb];
};

// acc[i] += sum_{k<32} T[i][k] * B[k * ldb + lane]   for 8 rows of the tile (row stride kNb)
__device__ __forceinline__ void mac_8x1(float (&acc)[8], const float* __restrict__ T, const float* __restrict__ B,
                                        const int ldb, const int lane)
{
#pragma unroll 2
    for (int k4 = 0; k4 < 32; k4 += 4) {
        float bv[4];
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) bv[kk] = B[(k4 + kk) * ldb + lane];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float4 a = *reinterpret_cast<const float4*>(T + i * kNb + k4);
            acc[i] = fmaf(a.x, bv[0], acc[i]); acc[i] = fmaf(a.y, bv[1], acc[i]);
            acc[i] = fmaf(a.z, bv[2], acc[i]); acc[i] = fmaf(a.w, bv[3], acc[i]);
        }
    }
}

__global__ void __launch_bounds__(256, 2)
lcol_kernel(const float* __restrict__ u11, int ldu, float* __restrict__ a21, int ld, int nrows,
            float* __restrict__ la_hi, float* __restrict__ la_lo, int ldla, int k_la)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LcolSmem& sm = *reinterpret_cast<LcolSmem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row0 = blockIdx.x * LH;

    // stage 0: the four 32x32 diagonal sub-blocks of U11 (needed first, for the inverses)
    for (int idx = tid; idx < 4 * 32 * 8; idx += 256) {
        const int bb = idx >> 8, r = 32 * bb + ((idx >> 3) & 31), c4 = 32 * bb + ((idx & 7) << 2);
        cp_async16(&sm.U[r][c4], u11 + (size_t)r * ldu + c4);
    }
    cp_async_commit();
    // stage 1: the rest of U11 and the 64 x 128 tile of rows
    for (int idx = tid; idx < kNb * kNb / 4; idx += 256) {
        const int r = idx >> 5, c4 = (idx & 31) << 2;
        if ((r >> 5) != (c4 >> 5)) cp_async16(&sm.U[r][c4], u11 + (size_t)r * ldu + c4);
    }
    for (int r = warp; r < LH; r += 8) {
        if (row0 + r < nrows) cp_async16(&sm.tile[r][4 * lane], a21 + (size_t)(row0 + r) * ld + 4 * lane);
        else *reinterpret_cast<float4*>(&sm.tile[r][4 * lane]) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    if (warp < 4) {
        // inverse of the 32x32 unit-upper diagonal sub-block `warp`: lane c solves U x = e_c from the
        // bottom up, the column held in registers
        const int o = 32 * warp;
        float x[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = (i == lane) ? 1.f : 0.f;
#pragma unroll
        for (int j = 31; j > 0; --j) {
#pragma unroll
            for (int i = 0; i < j; ++i) x[i] = fmaf(-sm.U[o + i][o + j], x[j], x[i]);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) sm.invUd[warp][i][lane] = x[i];
    }
    cp_async_wait<0>();
    __syncthreads();

    float* T = &sm.tile[8 * warp][0];                      // a warp only ever touches its own 8 rows of the tile
#pragma unroll 1
    for (int b = 0; b < 4; ++b) {
        // X_b = tile_b * inv(U_bb) : lane -> column lane of the sub-block
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
        mac_8x1(acc, T + 32 * b, &sm.invUd[b][0][0], 32, lane);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i) T[i * kNb + 32 * b + lane] = acc[i];
        __syncwarp();
        // columns right: tile[:, 32(b+1)+ ...] -= X_b * U[32b..32b+31, those columns], 32 columns per pass
#pragma unroll 1
        for (int cb = 32 * (b + 1); cb < kNb; cb += 32) {
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] = 0.f;
            mac_8x1(acc, T + 32 * b, &sm.U[32 * b][cb], kNb, lane);
#pragma unroll
            for (int i = 0; i < 8; ++i) T[i * kNb + cb + lane] -= acc[i];
        }
        __syncwarp();
    }
    __syncthreads();
    // write back: L in place, and (3xTF32 mode) its hi/lo split (A-role operand of the trailing update)
    for (int r = warp; r < LH; r += 8) {
        if (row0 + r >= nrows) break;
        const float4 v = *reinterpret_cast<const float4*>(&sm.tile[r][4 * lane]);
        *reinterpret_cast<float4*>(a21 + (size_t)(row0 + r) * ld + 4 * lane) = v;
        if (la_hi != nullptr) {
            float4 h, l;
            tf32_split2(v.x, h.x, l.x); tf32_split2(v.y, h.y, l.y); tf32_split2(v.z, h.z, l.z); tf32_split2(v.w, h.w, l.w);
            *reinterpret_cast<float4*>(la_hi + (size_t)(row0 + r) * ldla + k_la + 4 * lane) = h;
            *reinterpret_cast<float4*>(la_lo + (size_t)(row0 + r) * ldla + k_la + 4 * lane) = l;
        }
    }
}

void launch_lcol(const float* u11, int ldu, float* a21, int ld, int nrows,
                 float* la_hi, float* la_lo, int ldla, int k_la, cudaStream_t s)
{
    if (nrows <= 0) return;
    if (la_hi != nullptr && panel_mma_enabled()) launch_lcol_mma(u11, ldu, a21, ld, nrows, la_hi, la_lo, ldla, k_la, s);
    else lcol_kernel<<<(nrows + LH - 1) / LH, 256, sizeof(LcolSmem), s>>>(u11, ldu, a21, ld, nrows, la_hi, la_lo, ldla, k_la);
    ++g_launch_count;
}

// =====================================================================================
// 4. The two panel solves on the tensor cores (3xTF32 mode).
//    Both are the same computation: four 32-wide stages of  X_b = inv(D_bb) S_b  followed by
//    S_s -= O_sb X_b  for the later sub-blocks s > b, where the right-hand sides -- matrix columns for
//    the U-row solve, matrix rows for the L-col solve -- never interact.  A warp therefore owns 8 of them
//    for the whole solve and keeps all 128 x 8 values in mma accumulator registers; there is no
//    block-wide barrier between the stages.  The products run as warp-level m16n8k8 TF32 MMAs
//    (mma.sync: these are 32 x 32 x 8 products, far below a tcgen05 tile) with the same 3xTF32 split
//    as the trailing update -- lo*hi + hi*lo + hi*hi into an FP32 accumulator.  X_b goes through the
//    warp's own columns of the shared-memory strip to change from accumulator layout to B-operand layout.
//    Template strides: the constant operand element (m, k) is at A[m*SAM + k*SAK], the data element
//    (i, n) at D[i*SDI + n*SDN]  (i = index along the solve, n = which of the warp's 8 right-hand sides).
// =====================================================================================
// Operand split for the MMAs below: hi = x truncated to TF32 (the tensor core ignores the low 13 bits anyway),
// lo = x - hi (exact in FP32), truncated likewise.  Three ALU instructions; cvt.rna.tf32 is emulated in ~4 each
// on sm_100 and these kernels split every operand on the fly.  The representation error is <= 2^-21 |x|, the
// same order as the lo*lo term 3xTF32 drops.
template <uint32_t SIGN = 0u>                              // SIGN = 0x80000000: the split of -x
__device__ __forceinline__ void split_bits(const float x, uint32_t& hi, uint32_t& lo)
{
    const uint32_t h = __float_as_uint(x) & 0xffffe000u;
    hi = h ^ SIGN;
    lo = (__float_as_uint(x - __uint_as_float(h)) & 0xffffe000u) ^ SIGN;
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int SAM, int SAK, int SDI, int SDN, int BLK>
__device__ __forceinline__ void tri_blocks_mma(const float* __restrict__ inv, const float* __restrict__ off,
                                               float* __restrict__ D, const int lane)
{
    const int g = lane >> 2, t = lane & 3;
    float acc[8][4];
#pragma unroll
    for (int M = 0; M < 8; ++M) {
        acc[M][0] = D[(16 * M + g) * SDI + (2 * t) * SDN];
        acc[M][1] = D[(16 * M + g) * SDI + (2 * t + 1) * SDN];
        acc[M][2] = D[(16 * M + g + 8) * SDI + (2 * t) * SDN];
        acc[M][3] = D[(16 * M + g + 8) * SDI + (2 * t + 1) * SDN];
    }
    auto store_sub = [&](const int b) {                    // sub-block b of the warp's right-hand sides -> D
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            const int M = 2 * b + mt;
            D[(16 * M + g) * SDI + (2 * t) * SDN] = acc[M][0];
            D[(16 * M + g) * SDI + (2 * t + 1) * SDN] = acc[M][1];
            D[(16 * M + g + 8) * SDI + (2 * t) * SDN] = acc[M][2];
            D[(16 * M + g + 8) * SDI + (2 * t + 1) * SDN] = acc[M][3];
        }
    };
    auto load_b = [&](const int b, uint32_t (&bh)[4][2], uint32_t (&bl)[4][2]) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            split_bits(D[(32 * b + 8 * ks + t) * SDI + g * SDN], bh[ks][0], bl[ks][0]);
            split_bits(D[(32 * b + 8 * ks + t + 4) * SDI + g * SDN], bh[ks][1], bl[ks][1]);
        }
    };
    auto mac_tile = [&](float (&d)[4], const float* __restrict__ A, const int mt, auto neg,
                        const uint32_t (&bh)[4][2], const uint32_t (&bl)[4][2]) {
        constexpr uint32_t SG = decltype(neg)::value ? 0x80000000u : 0u;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            uint32_t ah[4], al[4];
            split_bits<SG>(A[(16 * mt + g) * SAM + (8 * ks + t) * SAK], ah[0], al[0]);
            split_bits<SG>(A[(16 * mt + g + 8) * SAM + (8 * ks + t) * SAK], ah[1], al[1]);
            split_bits<SG>(A[(16 * mt + g) * SAM + (8 * ks + t + 4) * SAK], ah[2], al[2]);
            split_bits<SG>(A[(16 * mt + g + 8) * SAM + (8 * ks + t + 4) * SAK], ah[3], al[3]);
            mma_tf32(d, al, bh[ks]);
            mma_tf32(d, ah, bl[ks]);
            mma_tf32(d, ah, bh[ks]);
        }
    };
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        uint32_t bh[4][2], bl[4][2];
        if (b > 0) { store_sub(b); __syncwarp(); }          // S_b with the updates of the earlier stages
        load_b(b, bh, bl);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {                     // X_b = inv(D_bb) S_b
            float d[4] = {0.f, 0.f, 0.f, 0.f};
            mac_tile(d, inv + b * BLK, mt, std::false_type{}, bh, bl);
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[2 * b + mt][e] = d[e];
        }
        __syncwarp();
        store_sub(b);
        __syncwarp();
        if (b < 3) {
            load_b(b, bh, bl);                               // X_b as the B operand
#pragma unroll
            for (int s_ = b + 1; s_ < 4; ++s_)               // S_s -= O_sb X_b
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
                    mac_tile(acc[2 * s_ + mt], off + (s_ * (s_ - 1) / 2 + b) * BLK, mt, std::true_type{}, bh, bl);
        }
    }
}

// (s, b) of the off-diagonal sub-block with id = s (s - 1) / 2 + b, s > b
__device__ __forceinline__ void offdiag_pair(const int id, int& s_, int& b)
{
    s_ = id < 1 ? 1 : (id < 3 ? 2 : 3);
    b = id - s_ * (s_ - 1) / 2;
}

constexpr int kUP = 36;                  // pitch of the 32 x 32 blocks of L11 (conflict-free A fragments: 4g + t)
constexpr int kUS = UW + 8;              // pitch of the strip (conflict-free B fragments: 8t + g)
struct UrowMmaSmem {
    float Ld[4][32][kUP];                // diagonal sub-blocks of L11
    float Lo[6][32][kUP];                // sub-blocks (s, b), s > b
    float inv[4][32][kUP];               // inverses of the diagonal sub-blocks
    float strip[kNb][kUS];
};

__device__ __forceinline__ void
urow_mma_body(unsigned char* smem_raw, const int bid, const float* __restrict__ l11, int ldl, float* __restrict__ a12, int ld, int ncols,
              float* __restrict__ ub_hi, float* __restrict__ ub_lo, int ldub, int k_ub)
{
    UrowMmaSmem& sm = *reinterpret_cast<UrowMmaSmem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c0 = bid * UW;

    for (int idx = tid; idx < 4 * 32 * 8; idx += 256) {
        const int bb = idx >> 8, r = (idx >> 3) & 31, c4 = (idx & 7) << 2;
        cp_async16(&sm.Ld[bb][r][c4], l11 + (size_t)(32 * bb + r) * ldl + 32 * bb + c4);
    }
    cp_async_commit();
    for (int idx = tid; idx < 6 * 32 * 8; idx += 256) {
        const int id = idx >> 8, r = (idx >> 3) & 31, c4 = (idx & 7) << 2;
        int s_, b;
        offdiag_pair(id, s_, b);
        cp_async16(&sm.Lo[id][r][c4], l11 + (size_t)(32 * s_ + r) * ldl + 32 * b + c4);
    }
    {
        const int c4 = (tid & 15) << 2;
        const bool ok = c0 + c4 < ncols;
        for (int r = tid >> 4; r < kNb; r += 16) {
            if (ok) cp_async16(&sm.strip[r][c4], a12 + (size_t)r * ld + c0 + c4);
            else *reinterpret_cast<float4*>(&sm.strip[r][c4]) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    if (warp < 4) {                                        // inverse of the lower-triangular sub-block `warp`
        const float myrinv = 1.0f / sm.Ld[warp][lane][lane];
        float x[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = (i == lane) ? 1.f : 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            x[j] *= __shfl_sync(0xffffffffu, myrinv, j);
#pragma unroll
            for (int i = j + 1; i < 32; ++i) x[i] = fmaf(-sm.Ld[warp][i][j], x[j], x[i]);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) sm.inv[warp][i][lane] = x[i];
    }
    cp_async_wait<0>();
    __syncthreads();

    tri_blocks_mma<kUP, 1, kUS, 1, 32 * kUP>(&sm.inv[0][0][0], &sm.Lo[0][0][0], &sm.strip[0][8 * warp], lane);
    __syncthreads();

    {
        const int c4 = (tid & 15) << 2;
        if (c0 + c4 < ncols)
            for (int r = tid >> 4; r < kNb; r += 16)
                *reinterpret_cast<float4*>(a12 + (size_t)r * ld + c0 + c4) = *reinterpret_cast<const float4*>(&sm.strip[r][c4]);
    }
    if (ub_hi != nullptr) {
        const int c = tid & 63, part = tid >> 6;           // 4 threads per column, 32 rows each
        if (c0 + c < ncols) {
            float* dh = ub_hi + (size_t)(c0 + c) * ldub + k_ub + 32 * part;
            float* dl = ub_lo + (size_t)(c0 + c) * ldub + k_ub + 32 * part;
#pragma unroll 2
            for (int r = 0; r < 32; r += 4) {
                float4 h, l;
                tf32_split2(sm.strip[32 * part + r][c], h.x, l.x);
                tf32_split2(sm.strip[32 * part + r + 1][c], h.y, l.y);
                tf32_split2(sm.strip[32 * part + r + 2][c], h.z, l.z);
                tf32_split2(sm.strip[32 * part + r + 3][c], h.w, l.w);
                *reinterpret_cast<float4*>(dh + r) = h;
                *reinterpret_cast<float4*>(dl + r) = l;
            }
        }
    }
}

constexpr int kLP = 40;                  // pitch of the 32 x 32 blocks of U11 (transposed A fragments: 8t + g)
constexpr int kLS = kNb + 4;             // pitch of the row tile (B fragments: 4g + t)
struct LcolMmaSmem {
    float Ud[4][32][kLP];                // diagonal sub-blocks of U11 (unit diagonal implied)
    float Uo[6][32][kLP];                // sub-blocks (rows of b, columns of s), s > b
    float inv[4][32][kLP];
    float tile[LH][kLS];
};

__device__ __forceinline__ void
lcol_mma_body(unsigned char* smem_raw, const int bid, const float* __restrict__ u11, int ldu, float* __restrict__ a21, int ld, int nrows,
              float* __restrict__ la_hi, float* __restrict__ la_lo, int ldla, int k_la)
{
    LcolMmaSmem& sm = *reinterpret_cast<LcolMmaSmem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row0 = bid * LH;

    for (int idx = tid; idx < 4 * 32 * 8; idx += 256) {
        const int bb = idx >> 8, r = (idx >> 3) & 31, c4 = (idx & 7) << 2;
        cp_async16(&sm.Ud[bb][r][c4], u11 + (size_t)(32 * bb + r) * ldu + 32 * bb + c4);
    }
    cp_async_commit();
    for (int idx = tid; idx < 6 * 32 * 8; idx += 256) {
        const int id = idx >> 8, r = (idx >> 3) & 31, c4 = (idx & 7) << 2;
        int s_, b;
        offdiag_pair(id, s_, b);
        cp_async16(&sm.Uo[id][r][c4], u11 + (size_t)(32 * b + r) * ldu + 32 * s_ + c4);
    }
    for (int r = warp; r < LH; r += 8) {
        if (row0 + r < nrows) cp_async16(&sm.tile[r][4 * lane], a21 + (size_t)(row0 + r) * ld + 4 * lane);
        else *reinterpret_cast<float4*>(&sm.tile[r][4 * lane]) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    if (warp < 4) {                                        // inverse of the unit-upper sub-block `warp`
        float x[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = (i == lane) ? 1.f : 0.f;
#pragma unroll
        for (int j = 31; j > 0; --j) {
#pragma unroll
            for (int i = 0; i < j; ++i) x[i] = fmaf(-sm.Ud[warp][i][j], x[j], x[i]);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) sm.inv[warp][i][lane] = x[i];
    }
    cp_async_wait<0>();
    __syncthreads();

    // transposed roles: X_b = S_b inv(U_bb) is  X_b^T = inv(U_bb)^T S_b^T, the constant operand is read transposed
    tri_blocks_mma<1, kLP, 1, kLS, 32 * kLP>(&sm.inv[0][0][0], &sm.Uo[0][0][0], &sm.tile[8 * warp][0], lane);
    __syncthreads();

    for (int r = warp; r < LH; r += 8) {
        if (row0 + r >= nrows) break;
        const float4 v = *reinterpret_cast<const float4*>(&sm.tile[r][4 * lane]);
        *reinterpret_cast<float4*>(a21 + (size_t)(row0 + r) * ld + 4 * lane) = v;
        if (la_hi != nullptr) {
            float4 h, l;
            tf32_split2(v.x, h.x, l.x); tf32_split2(v.y, h.y, l.y); tf32_split2(v.z, h.z, l.z); tf32_split2(v.w, h.w, l.w);
            *reinterpret_cast<float4*>(la_hi + (size_t)(row0 + r) * ldla + k_la + 4 * lane) = h;
            *reinterpret_cast<float4*>(la_lo + (size_t)(row0 + r) * ldla + k_la + 4 * lane) = l;
        }
    }
}

__global__ void __launch_bounds__(256, 2)
urow_mma_kernel(const float* __restrict__ l11, int ldl, float* __restrict__ a12, int ld, int ncols,
                float* __restrict__ ub_hi, float* __restrict__ ub_lo, int ldub, int k_ub)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    pdl_release();
    pdl_wait();
    urow_mma_body(smem_raw, blockIdx.x, l11, ldl, a12, ld, ncols, ub_hi, ub_lo, ldub, k_ub);
}
__global__ void __launch_bounds__(256, 2)
lcol_mma_kernel(const float* __restrict__ u11, int ldu, float* __restrict__ a21, int ld, int nrows,
                float* __restrict__ la_hi, float* __restrict__ la_lo, int ldla, int k_la)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    pdl_release();
    pdl_wait();
    lcol_mma_body(smem_raw, blockIdx.x, u11, ldu, a21, ld, nrows, la_hi, la_lo, ldla, k_la);
}
// Both solves of one inner block in one launch: they only depend on the diagonal factor, not on each other, and
// each of them is latency-bound (~10-13 us for any size below one wave), so sharing a launch takes one kernel off the
// panel chain.  The first n_ucta CTAs do the U-row columns, the rest the L-col rows.
__global__ void __launch_bounds__(256, 2)
urow_lcol_mma_kernel(const float* __restrict__ d11, int ldd, int n_ucta,
                     float* __restrict__ a12, int ld12, int ncols, float* __restrict__ ub_hi, float* __restrict__ ub_lo, int ldub, int k_ub,
                     float* __restrict__ a21, int ld21, int nrows, float* __restrict__ la_hi, float* __restrict__ la_lo, int ldla, int k_la)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    pdl_release();
    pdl_wait();
    if ((int)blockIdx.x < n_ucta) urow_mma_body(smem_raw, blockIdx.x, d11, ldd, a12, ld12, ncols, ub_hi, ub_lo, ldub, k_ub);
    else lcol_mma_body(smem_raw, (int)blockIdx.x - n_ucta, d11, ldd, a21, ld21, nrows, la_hi, la_lo, ldla, k_la);
}

void launch_urow_lcol(const float* d11, int ldd,
                      float* a12, int ld12, int ncols, float* ub_hi, float* ub_lo, int ldub, int k_ub,
                      float* a21, int ld21, int nrows, float* la_hi, float* la_lo, int ldla, int k_la, cudaStream_t s)
{
    if (ncols > 0 && nrows > 0 && ub_hi != nullptr && la_hi != nullptr && panel_mma_enabled() && panel_fuse_enabled()) {
        const int n_ucta = (ncols + UW - 1) / UW, n_lcta = (nrows + LH - 1) / LH;
        constexpr size_t smem = sizeof(UrowMmaSmem) > sizeof(LcolMmaSmem) ? sizeof(UrowMmaSmem) : sizeof(LcolMmaSmem);
        launch_k(urow_lcol_mma_kernel, dim3(n_ucta + n_lcta), dim3(256), smem, s, d11, ldd, n_ucta, a12, ld12, ncols, ub_hi, ub_lo, ldub, k_ub,
                 a21, ld21, nrows, la_hi, la_lo, ldla, k_la);
        ++g_launch_count;
        return;
    }
    launch_urow(d11, ldd, a12, ld12, ncols, ub_hi, ub_lo, ldub, k_ub, s);
    launch_lcol(d11, ldd, a21, ld21, nrows, la_hi, la_lo, ldla, k_la, s);
}

static void launch_urow_mma(const float* l11, int ldl, float* a12, int ld, int ncols, float* ub_hi, float* ub_lo, int ldub, int k_ub, cudaStream_t s)
{
    launch_k(urow_mma_kernel, dim3((ncols + UW - 1) / UW), dim3(256), sizeof(UrowMmaSmem), s, l11, ldl, a12, ld, ncols, ub_hi, ub_lo, ldub, k_ub);
}
static void launch_lcol_mma(const float* u11, int ldu, float* a21, int ld, int nrows, float* la_hi, float* la_lo, int ldla, int k_la, cudaStream_t s)
{
    launch_k(lcol_mma_kernel, dim3((nrows + LH - 1) / LH), dim3(256), sizeof(LcolMmaSmem), s, u11, ldu, a21, ld, nrows, la_hi, la_lo, ldla, k_la);
}

void panel_setup_device()
{
    cudaFuncSetAttribute(urow_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UrowMmaSmem));
    cudaFuncSetAttribute(urow_lcol_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)(sizeof(UrowMmaSmem) > sizeof(LcolMmaSmem) ? sizeof(UrowMmaSmem) : sizeof(LcolMmaSmem)));
    cudaFuncSetAttribute(lcol_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LcolMmaSmem));
    cudaFuncSetAttribute(urow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UrowSmem));
    cudaFuncSetAttribute(lcol_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LcolSmem));
}

} // namespace gb200
