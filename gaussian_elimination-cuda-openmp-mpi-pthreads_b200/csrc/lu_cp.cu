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
//
// Fused with their collective: the two kernels that produce the "crit" package (ext_finish: inv(U11)^T split; crit_split:
// U[K, K+1]^T split) store it straight into the NEXT owner's receive buffer as well -- plain st.global on a peer mapping
// over NVLink, no TMA involved -- and the last CTA to finish publishes a sequence number in the peer's flag word
// (fence.sys + st.release.sys).  The receiver's stream waits on that word.  This replaces, on the critical path, the
// reference's per-step MPI_Bcast of the pivot row (gauss_mpi/gauss_internal_input.c:135-137) for the one rank that needs
// the data first; everyone else still gets it by the NCCL broadcast, with a step of slack.
#include "gauss_internal.cuh"
#include <stdio.h>

namespace gb200 {

// Called by every thread of every CTA at the end of a kernel that wrote to peer memory: the last CTA to arrive stores
// `value` into `flag` (a word in the peer's memory) after all CTAs' stores have been made visible system-wide.
__device__ __forceinline__ void signal_peer_when_done(unsigned* counter, unsigned nblocks, unsigned* flag, unsigned value)
{
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        __threadfence_system();
        const unsigned prev = atomicAdd(counter, 1u);
        if (prev == nblocks - 1) {
            *counter = 0;                                     // rearm for the next launch (stream order makes it visible)
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
        }
    }
}

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
    pdl_release();
    pdl_wait();
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
    launch_k(ext_prepare_kernel, grid, dim3(256), 0, s, a11, ld, E, lde, kb);
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
                  float* __restrict__ il_hi, float* __restrict__ il_lo, float* __restrict__ iut_hi, float* __restrict__ iut_lo, int ldo,
                  float* __restrict__ peer_hi, float* __restrict__ peer_lo, unsigned* counter, unsigned* peer_flag, unsigned seq,
                  float* __restrict__ iu_raw)
{
    __shared__ float t[32][33];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int c0 = blockIdx.x * 32;
    const int nby = kb / 32;
    pdl_release();
    pdl_wait();
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
        for (int i = ty; i < 32; i += 8) {
            const float v = E[(size_t)(kb + r0 + i) * lde + c0 + tx];
            t[i][tx] = v;
            if (iu_raw != nullptr) iu_raw[(size_t)(r0 + i) * ldo + c0 + tx] = v;     // inv(U11) itself, kept for back substitution
        }
        __syncthreads();
        for (int i = ty; i < 32; i += 8) {
            float h, l;
            cp_split(t[tx][i], h, l);                            // element (row r0+tx, column c0+i) -> out[c0+i][r0+tx]
            iut_hi[(size_t)(c0 + i) * ldo + r0 + tx] = h;
            iut_lo[(size_t)(c0 + i) * ldo + r0 + tx] = l;
            if (peer_hi != nullptr) {                            // the next owner's copy, over NVLink
                peer_hi[(size_t)(c0 + i) * ldo + r0 + tx] = h;
                peer_lo[(size_t)(c0 + i) * ldo + r0 + tx] = l;
            }
        }
    }
    if (peer_flag != nullptr) signal_peer_when_done(counter, gridDim.x * gridDim.y, peer_flag, seq);
}

void launch_ext_finish(const float* E, int lde, int kb, float* a11, int ld,
                       float* il_hi, float* il_lo, float* iut_hi, float* iut_lo, int ldo,
                       float* peer_hi, float* peer_lo, unsigned* counter, unsigned* peer_flag, unsigned seq, float* iu_raw, cudaStream_t s)
{
    dim3 grid(kb / 32, 2 * (kb / 32));
    launch_k(ext_finish_kernel, grid, dim3(32, 8), 0, s, E, lde, kb, a11, ld, il_hi, il_lo, iut_hi, iut_lo, ldo, peer_hi, peer_lo, counter, peer_flag, seq, iu_raw);
    ++g_launch_count;
}

// hi/lo[r][c] = split(src[r][c]) for r < rows, c < cols (cols a multiple of 4), written to the local operand buffers and,
// when given, to the next owner's; then the peer's flag word is set to `seq`.
__global__ void __launch_bounds__(128)
crit_split_kernel(const float* __restrict__ src, int lds, float* __restrict__ hi, float* __restrict__ lo, int ldd, int rows, int cols,
                  float* __restrict__ peer_hi, float* __restrict__ peer_lo, unsigned* counter, unsigned* peer_flag, unsigned seq)
{
    const int c4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int r = blockIdx.y;
    pdl_release();
    pdl_wait();
    if (c4 < cols && r < rows) {
        const float4 v = *reinterpret_cast<const float4*>(src + (size_t)r * lds + c4);
        float4 h, l;
        cp_split(v.x, h.x, l.x); cp_split(v.y, h.y, l.y); cp_split(v.z, h.z, l.z); cp_split(v.w, h.w, l.w);
        *reinterpret_cast<float4*>(hi + (size_t)r * ldd + c4) = h;
        *reinterpret_cast<float4*>(lo + (size_t)r * ldd + c4) = l;
        if (peer_hi != nullptr) {
            *reinterpret_cast<float4*>(peer_hi + (size_t)r * ldd + c4) = h;
            *reinterpret_cast<float4*>(peer_lo + (size_t)r * ldd + c4) = l;
        }
    }
    if (peer_flag != nullptr) signal_peer_when_done(counter, gridDim.x * gridDim.y, peer_flag, seq);
}

void launch_crit_split(const float* src, int lds, float* hi, float* lo, int ldd, int rows, int cols,
                       float* peer_hi, float* peer_lo, unsigned* counter, unsigned* peer_flag, unsigned seq, cudaStream_t s)
{
    dim3 grid((cols / 4 + 127) / 128, rows);
    launch_k(crit_split_kernel, grid, dim3(128), 0, s, src, lds, hi, lo, ldd, rows, cols, peer_hi, peer_lo, counter, peer_flag, seq);
    ++g_launch_count;
}

// Back substitution step of one outer block on its owner, fused with its broadcast (solveGauss,
// gauss_internal_input.c:144-159, for the rows of the block):  x_K = inv(U11) * y_K  as a matrix-vector product (a warp
// per row; y_K is the right-hand-side column of the block's rows, already reduced by the blocks solved before), stored
// into the owner's solution vector and straight into every peer's receive vector over NVLink; the last CTA then sets
// every peer's flag word.  Replaces, per outer block, a chained triangular solve plus one NCCL broadcast.
__global__ void __launch_bounds__(256)
bs_solve_push_kernel(const float* __restrict__ iu, int ldiu, const float* __restrict__ ycol, int ld, int kb,
                     float* __restrict__ x_local, BsPeers peers, unsigned* counter, unsigned seq)
{
    extern __shared__ float ysm[];
    for (int j = threadIdx.x; j < kb; j += blockDim.x) ysm[j] = ycol[(size_t)j * ld];
    __syncthreads();
    const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row < kb) {
        float acc = 0.f;
        const int j0 = (row / 128) * 128;                         // inv(U11) is upper triangular: columns left of the row's 128-block are zero
        for (int j = j0 + 4 * lane; j < kb; j += 128) {
            const float4 a = *reinterpret_cast<const float4*>(iu + (size_t)row * ldiu + j);
            acc = fmaf(a.x, ysm[j], acc); acc = fmaf(a.y, ysm[j + 1], acc); acc = fmaf(a.z, ysm[j + 2], acc); acc = fmaf(a.w, ysm[j + 3], acc);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (lane == 0) {
            x_local[row] = acc;
            for (int p = 0; p < peers.n; ++p) peers.x[p][row] = acc;
        }
    }
    if (peers.n > 0) {
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            if (atomicAdd(counter, 1u) == gridDim.x - 1) {
                *counter = 0;
                __threadfence_system();
                for (int p = 0; p < peers.n; ++p)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peers.flag[p]), "r"(seq) : "memory");
            }
        }
    }
}
void launch_bs_solve_push(const float* iu, int ldiu, const float* ycol, int ld, int kb, float* x_local, const BsPeers& peers,
                          unsigned* counter, unsigned seq, cudaStream_t s)
{
    bs_solve_push_kernel<<<(kb + 7) / 8, 256, kb * sizeof(float), s>>>(iu, ldiu, ycol, ld, kb, x_local, peers, counter, seq);
    ++g_launch_count;
}

// Fallback for drivers without stream memory operations: one warp spins (bounded) until *flag >= value.
__global__ void wait_flag_kernel(const unsigned* flag, unsigned value)
{
    if (threadIdx.x != 0) return;
    const long long t0 = clock64();
    for (;;) {
        unsigned v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if ((int)(v - value) >= 0) break;
        if (clock64() - t0 > 20000000000LL) { printf("gauss_cuda: peer flag wait timed out (have %u, want %u)\n", v, value); __trap(); }
        __nanosleep(100);
    }
}
void launch_wait_flag(const unsigned* flag, unsigned value, cudaStream_t s)
{
    wait_flag_kernel<<<1, 32, 0, s>>>(flag, value);
    ++g_launch_count;
}

} // namespace gb200
