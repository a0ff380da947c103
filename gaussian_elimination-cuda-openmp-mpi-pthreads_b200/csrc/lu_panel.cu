// Panel kernels of the blocked no-pivot LU (the latency-critical part of computeGauss,
// gauss_internal_input.c:118-138): the 128x128 diagonal-block factor and the two triangular
// panel solves.  All three work on 32-wide sub-blocks so that the only strictly sequential part
// is 32 pivot steps at a time; everything else is small dense block products spread over 256 threads.
//
//   diag_factor  A11 = L11*U11 in shared memory (Crout: U unit upper, L carries the pivots;
//                getPivot :100-111 scales row i by 1/pivot, computeGauss :129-136 eliminates)
//   urow         U12 = L11^-1 * A12      (the part of rows i right of the diagonal block, :132-134)
//   lcol         L21 = A21 * U11^-1      (the multipliers a_ji of the rows below, :130)
// urow/lcol invert the four 32x32 diagonal sub-blocks once per CTA (one warp each) and then run
// block forward substitution:  X_b = inv(D_b) * (A_b - sum_{c<b} L_bc X_c)  as dense products.
#include "gauss_internal.cuh"

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
__device__ __forceinline__ void bar_named(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// =====================================================================================
// 1. Diagonal block factor
// =====================================================================================
constexpr int DLD = kNb + 4;      // shared-memory row stride (floats)
struct DiagSmem {
    float S[kNb][DLD];
    float rowbuf[2][32];
    float rinv[kNb];
    int first_bad;
};

template <int B>
__device__ __forceinline__ void diag_block(DiagSmem& sm, const int tid)
{
    constexpr int o = 32 * B;          // first row/column of this sub-block
    constexpr int R = kNb - o;         // rows in the column panel
    constexpr int Cb = R - 32;         // columns right of the panel inside the 128 block
    // (a) factor the (R x 32) column panel: thread t owns row o+t, 32 entries in registers; one
    //     named barrier per pivot step, the normalised pivot row travels through rowbuf
    if (tid < R) {
        float p[32];
        const float* src = &sm.S[o + tid][o];
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const float4 v = *reinterpret_cast<const float4*>(src + j);
            p[j] = v.x; p[j + 1] = v.y; p[j + 2] = v.z; p[j + 3] = v.w;
        }
#pragma unroll
        for (int kk = 0; kk < 32; ++kk) {
            float* rb = sm.rowbuf[kk & 1];
            if (tid == kk) {
                const float piv = p[kk];
                if (!(fabsf(piv) > 0.f) || isinf(piv)) atomicMin(&sm.first_bad, o + kk);
                const float r = 1.0f / piv;                    // getPivot :105-107 (reciprocal instead of 31 divisions)
                sm.rinv[o + kk] = r;
#pragma unroll
                for (int j = kk + 1; j < 32; ++j) { p[j] *= r; rb[j] = p[j]; }
            }
            bar_named(1, R);
            if (tid > kk) {
                const float l = p[kk];
#pragma unroll
                for (int j = kk + 1; j < 32; ++j) p[j] = fmaf(-l, rb[j], p[j]);     // computeGauss :132-134
            }
        }
        float* dst = &sm.S[o + tid][o];
#pragma unroll
        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(p[j], p[j + 1], p[j + 2], p[j + 3]);
    }
    __syncthreads();
    if (Cb > 0) {
        // (b) block row of U right of the panel: one thread per column, forward substitution with the 32x32 L
        if (tid < Cb) {
            const int c = o + 32 + tid;
            float x[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) x[i] = sm.S[o + i][c];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                x[j] *= sm.rinv[o + j];
#pragma unroll
                for (int i = j + 1; i < 32; ++i) x[i] = fmaf(-sm.S[o + i][o + j], x[j], x[i]);
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) sm.S[o + i][c] = x[i];
        }
        __syncthreads();
        // (c) rank-32 update of the trailing (Cb x Cb) part, cyclic T x T register tiles
        constexpr int T = Cb / 16 > 0 ? Cb / 16 : 1;
        const int ty = tid >> 4, tx = tid & 15;
        float acc[T][T];
#pragma unroll
        for (int i = 0; i < T; ++i)
#pragma unroll
            for (int j = 0; j < T; ++j) acc[i][j] = 0.f;
#pragma unroll 8
        for (int k = 0; k < 32; ++k) {
            float a[T], bv[T];
#pragma unroll
            for (int i = 0; i < T; ++i) a[i] = sm.S[o + 32 + ty + 16 * i][o + k];
#pragma unroll
            for (int j = 0; j < T; ++j) bv[j] = sm.S[o + k][o + 32 + tx + 16 * j];
#pragma unroll
            for (int i = 0; i < T; ++i)
#pragma unroll
                for (int j = 0; j < T; ++j) acc[i][j] = fmaf(a[i], bv[j], acc[i][j]);
        }
#pragma unroll
        for (int i = 0; i < T; ++i)
#pragma unroll
            for (int j = 0; j < T; ++j) sm.S[o + 32 + ty + 16 * i][o + 32 + tx + 16 * j] -= acc[i][j];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256, 1)
diag_factor_kernel(float* __restrict__ blk, int ld, int grow0, int* __restrict__ status)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    DiagSmem& sm = *reinterpret_cast<DiagSmem*>(smem_raw);
    const int tid = threadIdx.x;
    if (tid == 0) sm.first_bad = kNb;
    for (int idx = tid; idx < kNb * kNb / 4; idx += 256) {
        const int r = idx >> 5, c4 = (idx & 31) << 2;
        *reinterpret_cast<float4*>(&sm.S[r][c4]) = *reinterpret_cast<const float4*>(blk + (size_t)r * ld + c4);
    }
    __syncthreads();
    diag_block<0>(sm, tid);
    diag_block<1>(sm, tid);
    diag_block<2>(sm, tid);
    diag_block<3>(sm, tid);
    for (int idx = tid; idx < kNb * kNb / 4; idx += 256) {
        const int r = idx >> 5, c4 = (idx & 31) << 2;
        *reinterpret_cast<float4*>(blk + (size_t)r * ld + c4) = *reinterpret_cast<const float4*>(&sm.S[r][c4]);
    }
    // earliest bad pivot of this block; the first block to report wins (blocks run in pivot order)
    if (tid == 0 && sm.first_bad < kNb) atomicCAS(status, 0, grow0 + sm.first_bad + 1);
}

void launch_diag_factor(float* blk, int ld, int global_row0, int* status, cudaStream_t s)
{
    diag_factor_kernel<<<1, 256, sizeof(DiagSmem), s>>>(blk, ld, global_row0, status);
    ++g_launch_count;
}

// =====================================================================================
// 2. U-row panel:  X = L11^-1 * A12, 64 matrix columns per CTA, 256 threads
// =====================================================================================
constexpr int UW = 64;
struct UrowSmem {
    float L[kNb][kNb];            // row-major L11 (i >= j part used)
    float invLd[4][32][32];       // inverses of the four 32x32 diagonal sub-blocks of L11, row-major
    float strip[kNb][UW];
};

// rows [row0, row0 + 8*RPW) of the strip  -=  L[rows, 32b..32b+31] * strip[32b..32b+31, :]
template <int RPW>
__device__ __forceinline__ void urow_update(UrowSmem& sm, const int b, const int row0, const int warp, const int lane)
{
    float acc[RPW][2];
#pragma unroll
    for (int i = 0; i < RPW; ++i) { acc[i][0] = 0.f; acc[i][1] = 0.f; }
    const int r0 = row0 + RPW * warp;
#pragma unroll
    for (int k4 = 0; k4 < 32; k4 += 4) {
        float4 a[RPW];
#pragma unroll
        for (int i = 0; i < RPW; ++i) a[i] = *reinterpret_cast<const float4*>(&sm.L[r0 + i][32 * b + k4]);
        float bv[4][2];
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) { bv[kk][0] = sm.strip[32 * b + k4 + kk][lane]; bv[kk][1] = sm.strip[32 * b + k4 + kk][lane + 32]; }
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
            acc[i][0] = fmaf(a[i].x, bv[0][0], acc[i][0]); acc[i][1] = fmaf(a[i].x, bv[0][1], acc[i][1]);
            acc[i][0] = fmaf(a[i].y, bv[1][0], acc[i][0]); acc[i][1] = fmaf(a[i].y, bv[1][1], acc[i][1]);
            acc[i][0] = fmaf(a[i].z, bv[2][0], acc[i][0]); acc[i][1] = fmaf(a[i].z, bv[2][1], acc[i][1]);
            acc[i][0] = fmaf(a[i].w, bv[3][0], acc[i][0]); acc[i][1] = fmaf(a[i].w, bv[3][1], acc[i][1]);
        }
    }
#pragma unroll
    for (int i = 0; i < RPW; ++i) { sm.strip[r0 + i][lane] -= acc[i][0]; sm.strip[r0 + i][lane + 32] -= acc[i][1]; }
}

__global__ void __launch_bounds__(256, 2)
urow_kernel(const float* __restrict__ l11, int ldl, float* __restrict__ a12, int ld, int ncols,
            float* __restrict__ ub_hi, float* __restrict__ ub_lo, int ldub, int k_ub)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    UrowSmem& sm = *reinterpret_cast<UrowSmem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c0 = blockIdx.x * UW;

    for (int idx = tid; idx < kNb * kNb / 4; idx += 256) {
        const int r = idx >> 5, c4 = (idx & 31) << 2;
        *reinterpret_cast<float4*>(&sm.L[r][c4]) = *reinterpret_cast<const float4*>(l11 + (size_t)r * ldl + c4);
    }
    // strip: 128 rows x 64 columns, float4 per thread (columns beyond ncols are the zero padding of W)
    {
        const int c4 = (tid & 15) << 2;
        const bool ok = c0 + c4 < ncols;
        for (int r = tid >> 4; r < kNb; r += 16) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok) v = *reinterpret_cast<const float4*>(a12 + (size_t)r * ld + c0 + c4);
            *reinterpret_cast<float4*>(&sm.strip[r][c4]) = v;
        }
    }
    __syncthreads();
    if (warp < 4) {
        // inverse of the 32x32 lower-triangular diagonal sub-block `warp`: lane c solves L x = e_c
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
    __syncthreads();

#pragma unroll 1
    for (int b = 0; b < 4; ++b) {
        // X_b = inv(L_bb) * strip_b : warp w -> rows 4w..4w+3 of the sub-block, lane -> columns lane, lane+32
        float acc[4][2];
#pragma unroll
        for (int i = 0; i < 4; ++i) { acc[i][0] = 0.f; acc[i][1] = 0.f; }
#pragma unroll
        for (int k4 = 0; k4 < 32; k4 += 4) {
            float4 a[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(&sm.invLd[b][4 * warp + i][k4]);
            float bv[4][2];
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) { bv[kk][0] = sm.strip[32 * b + k4 + kk][lane]; bv[kk][1] = sm.strip[32 * b + k4 + kk][lane + 32]; }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                acc[i][0] = fmaf(a[i].x, bv[0][0], acc[i][0]); acc[i][1] = fmaf(a[i].x, bv[0][1], acc[i][1]);
                acc[i][0] = fmaf(a[i].y, bv[1][0], acc[i][0]); acc[i][1] = fmaf(a[i].y, bv[1][1], acc[i][1]);
                acc[i][0] = fmaf(a[i].z, bv[2][0], acc[i][0]); acc[i][1] = fmaf(a[i].z, bv[2][1], acc[i][1]);
                acc[i][0] = fmaf(a[i].w, bv[3][0], acc[i][0]); acc[i][1] = fmaf(a[i].w, bv[3][1], acc[i][1]);
            }
        }
        __syncthreads();                                   // every read of strip_b is done
#pragma unroll
        for (int i = 0; i < 4; ++i) { sm.strip[32 * b + 4 * warp + i][lane] = acc[i][0]; sm.strip[32 * b + 4 * warp + i][lane + 32] = acc[i][1]; }
        __syncthreads();
        if (b == 0)      urow_update<12>(sm, 0, 32, warp, lane);
        else if (b == 1) urow_update<8>(sm, 1, 64, warp, lane);
        else if (b == 2) urow_update<4>(sm, 2, 96, warp, lane);
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
#pragma unroll
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

void launch_urow(const float* l11, int ldl, float* a12, int ld, int ncols,
                 float* ub_hi, float* ub_lo, int ldub, int k_ub, cudaStream_t s)
{
    if (ncols <= 0) return;
    urow_kernel<<<(ncols + UW - 1) / UW, 256, sizeof(UrowSmem), s>>>(l11, ldl, a12, ld, ncols, ub_hi, ub_lo, ldub, k_ub);
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

// tile[:, 32(b+1) ...] -= tile[:, 32b..32b+31] * U[32b..32b+31, 32(b+1) ...]   (NJ 32-column groups)
template <int NJ>
__device__ __forceinline__ void lcol_update(LcolSmem& sm, const int b, const int warp, const int lane)
{
    float acc[8][NJ];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j] = 0.f;
    const int r0 = 8 * warp, cbase = 32 * (b + 1) + lane;
#pragma unroll
    for (int k4 = 0; k4 < 32; k4 += 4) {
        float4 a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4*>(&sm.tile[r0 + i][32 * b + k4]);
        float bv[4][NJ];
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
#pragma unroll
            for (int j = 0; j < NJ; ++j) bv[kk][j] = sm.U[32 * b + k4 + kk][cbase + 32 * j];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                acc[i][j] = fmaf(a[i].x, bv[0][j], acc[i][j]);
                acc[i][j] = fmaf(a[i].y, bv[1][j], acc[i][j]);
                acc[i][j] = fmaf(a[i].z, bv[2][j], acc[i][j]);
                acc[i][j] = fmaf(a[i].w, bv[3][j], acc[i][j]);
            }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) sm.tile[r0 + i][cbase + 32 * j] -= acc[i][j];
}

__global__ void __launch_bounds__(256, 2)
lcol_kernel(const float* __restrict__ u11, int ldu, float* __restrict__ a21, int ld, int nrows,
            float* __restrict__ la_hi, float* __restrict__ la_lo, int ldla, int k_la)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LcolSmem& sm = *reinterpret_cast<LcolSmem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row0 = blockIdx.x * LH;

    for (int idx = tid; idx < kNb * kNb / 4; idx += 256) {
        const int r = idx >> 5, c4 = (idx & 31) << 2;
        *reinterpret_cast<float4*>(&sm.U[r][c4]) = *reinterpret_cast<const float4*>(u11 + (size_t)r * ldu + c4);
    }
    for (int r = warp; r < LH; r += 8) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < nrows) v = *reinterpret_cast<const float4*>(a21 + (size_t)(row0 + r) * ld + 4 * lane);
        *reinterpret_cast<float4*>(&sm.tile[r][4 * lane]) = v;
    }
    __syncthreads();
    if (warp < 4) {
        // inverse of the 32x32 unit-upper diagonal sub-block `warp`: lane c solves U x = e_c from the bottom up
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
    __syncthreads();

#pragma unroll 1
    for (int b = 0; b < 4; ++b) {
        // X_b = tile_b * inv(U_bb) : warp w -> rows 8w..8w+7, lane -> column lane of the sub-block
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
        for (int k4 = 0; k4 < 32; k4 += 4) {
            float bv[4];
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) bv[kk] = sm.invUd[b][k4 + kk][lane];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 a = *reinterpret_cast<const float4*>(&sm.tile[8 * warp + i][32 * b + k4]);
                acc[i] = fmaf(a.x, bv[0], acc[i]); acc[i] = fmaf(a.y, bv[1], acc[i]);
                acc[i] = fmaf(a.z, bv[2], acc[i]); acc[i] = fmaf(a.w, bv[3], acc[i]);
            }
        }
        __syncwarp();                                       // a warp only reads/writes its own 8 rows of tile_b
#pragma unroll
        for (int i = 0; i < 8; ++i) sm.tile[8 * warp + i][32 * b + lane] = acc[i];
        __syncwarp();
        if (b == 0)      lcol_update<3>(sm, 0, warp, lane);
        else if (b == 1) lcol_update<2>(sm, 1, warp, lane);
        else if (b == 2) lcol_update<1>(sm, 2, warp, lane);
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
    lcol_kernel<<<(nrows + LH - 1) / LH, 256, sizeof(LcolSmem), s>>>(u11, ldu, a21, ld, nrows, la_hi, la_lo, ldla, k_la);
    ++g_launch_count;
}

void panel_setup_device()
{
    cudaFuncSetAttribute(diag_factor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(DiagSmem));
    cudaFuncSetAttribute(urow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UrowSmem));
    cudaFuncSetAttribute(lcol_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LcolSmem));
}

} // namespace gb200
