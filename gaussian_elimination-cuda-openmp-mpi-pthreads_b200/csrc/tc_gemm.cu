// Tensor-core tile engine for sm_100a: C (+)= -/+ (Ahi + Alo) * (Bhi + Blo)^T as split 3xTF32.
//
// One persistent CTA per SM, 192 threads, warp-specialised:
//   warp 0      TMA producer   cp.async.bulk.tensor loads of the four operand tiles of a k-slice
//                              (A hi/lo: 128 x BK, B hi/lo: 256 x BK, K-major, hardware swizzle)
//   warp 1      MMA issuer     one elected lane issues tcgen05.mma.kind::tf32, three per 8-wide k step
//                              (lo*hi, hi*lo, hi*hi - small terms first) into a 128x256 FP32 accumulator
//                              in TMEM; accumulators are double buffered (2 x 256 columns = all of TMEM)
//   warps 2..5  epilogue       tcgen05.ld 32 columns at a time -> swizzled smem staging -> TMA
//                              reduce-add into C (trailing update  A22 -= L21*U12; the MMA negates A)
//                              or TMA store (matmul, beta = 0)
// smem ring and TMEM hand-offs use mbarriers; every wait is bounded and traps instead of hanging.
//
// It serves both the trailing update of the blocked LU (computeGauss' inner k loop,
// gauss_internal_input.c:132-134, as a rank-kb update) and matmul_cuda (cuda_matmul.cu:89-101).
#include "gauss_internal.cuh"
#include "gauss_cuda.h"

#include <cuda.h>
#include <stdio.h>
#include <algorithm>
#include <atomic>

namespace gb200 {

constexpr int TBM = 128;          // tile rows   (UMMA M, cta_group::1)
constexpr int TBN = 256;          // tile cols   (UMMA N)
constexpr int TEPI_W = 32;        // epilogue chunk: 32 fp32 columns = one 128 B swizzle row
constexpr int TSTAGE_C = TBM * TEPI_W * 4;   // 16 KB staging buffer
constexpr int kTcThreads = 192;

template <int BK, int STAGES>
struct TcCfg {
    static constexpr int a_bytes = TBM * BK * 4;
    static constexpr int b_bytes = TBN * BK * 4;
    static constexpr int stage_bytes = 2 * a_bytes + 2 * b_bytes;
    static constexpr int ring_bytes = STAGES * stage_bytes;
    static constexpr int bar_off = ring_bytes + 2 * TSTAGE_C;
    static constexpr int smem_bytes = bar_off + 256 + 1024;     // barriers + slack for 1024 B alignment
    static constexpr int swizzle_bytes = BK * 4;                // 64 or 128
};

// ---------------------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug must surface as a trapped kernel, never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) {
            printf("tc_gemm: mbarrier wait timed out (block %d thread %d bar 0x%x parity %u)\n", blockIdx.x, threadIdx.x, bar, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .b32 r;\n\t.reg .pred p;\n\t"
        "elect.sync r|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint32_t bar, uint32_t dst, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1)
{
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* map, uint32_t src, int c0, int c1)
{
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void tma_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile in shared memory, rows of SWZ bytes, 8-row swizzle atoms (SBO = 8 * SWZ).
template <int SWZ>
__device__ __forceinline__ uint64_t make_kmajor_desc(uint32_t saddr)
{
    constexpr uint64_t layout = SWZ == 128 ? 2 : (SWZ == 64 ? 4 : 6);
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);               // start address        bits [0,14)
    d |= (uint64_t)1 << 16;                               // leading byte offset  bits [16,30)  (unused for swizzled K-major)
    d |= (uint64_t)((8 * SWZ) >> 4) << 32;                // stride byte offset   bits [32,46)
    d |= (uint64_t)1 << 46;                               // descriptor version   bits [46,48)  (Blackwell)
    d |= layout << 61;                                    // swizzle mode         bits [61,64)
    return d;
}

// Tile rasterisation.  Tiles are walked in bands of kBandN tile-columns, row-major inside a band, so
// the ~148 tiles in flight cover a compact (about 9 x 16) patch: the band's B slab (16 x 256 rows of
// split U, 16 MB at K = 512) and the patch's A tiles stay L2-resident while C streams through.
// (Plain row-major order re-streamed all of B for every tile row: 3.4x read amplification at N = 32768.)
constexpr int kBandN = 16;
__device__ __forceinline__ void tile_coords(int t, int num_m, int num_n, int& m0, int& n0)
{
    const int band_tiles = num_m * kBandN;
    const int band = t / band_tiles;
    const int in_band = t - band * band_tiles;
    const int wb = min(kBandN, num_n - band * kBandN);      // the last band may be narrower
    const int mi = in_band / wb;
    const int ni = band * kBandN + (in_band - mi * wb);
    m0 = mi * TBM;
    n0 = ni * TBN;
}

// ---------------------------------------------------------------------------- the kernel
template <int BK, int STAGES>
__global__ void __launch_bounds__(kTcThreads, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tm_ahi, const __grid_constant__ CUtensorMap tm_alo,
               const __grid_constant__ CUtensorMap tm_bhi, const __grid_constant__ CUtensorMap tm_blo,
               const __grid_constant__ CUtensorMap tm_c,
               int m, int n, int k, int negate, int reduce, int* __restrict__ sched, int dynamic, int max_tiles)
{
    using Cfg = TcCfg<BK, STAGES>;
    extern __shared__ unsigned char smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;      // swizzle atoms need 1024 B alignment
    const uint32_t bar_base = base + Cfg::bar_off;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    // dynamic tile scheduler: a 4-deep queue of tile indices, producer -> {MMA issuer, epilogue}
    auto qfull_bar = [&](int q) { return bar_base + 8u * (2 * STAGES + 5 + q); };
    auto qempty_bar = [&](int q) { return bar_base + 8u * (2 * STAGES + 9 + q); };
    unsigned char* gen_base = smem_dyn + (base - smem_u32(smem_dyn));
    volatile int* q_tile = reinterpret_cast<volatile int*>(gen_base + Cfg::bar_off + 8 * (2 * STAGES + 13));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    pdl_release();
    const int num_n = (n + TBN - 1) / TBN;
    const int num_m = (m + TBM - 1) / TBM;
    const int num_tiles = num_m * num_n;
    const int num_kb = k / BK;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 128); }
        for (int q = 0; q < 4; ++q) { mbar_init(qfull_bar(q), 1); mbar_init(qempty_bar(q), 1 + 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(gen_base + Cfg::bar_off + 8 * (2 * STAGES + 4));
    pdl_wait();                                             // predecessors' operands / C are complete from here on

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            int stage = 0; uint32_t phase = 0;
            // first tile is static (blockIdx.x); later ones come from a global counter, so a CTA that
            // gets its SM late (or shares the GPU with another stream's kernels) simply takes fewer tiles
            for (int it = 0;; ++it) {
                const int q = it & 3;
                // dynamic == 1: first tile static, the rest from the counter; dynamic == 2 (bounded CTAs, grid larger
                // than the SM count): every tile from the counter so the tile order stays the band order
                int t;
                if (max_tiles > 0 && it >= max_tiles) t = -1;                       // this CTA has had its share
                else if (dynamic == 2) t = atomicAdd(sched, 1);
                else if (it == 0 || !dynamic) t = (int)blockIdx.x + it * (int)gridDim.x;
                else t = atomicAdd(sched, 1) + (int)gridDim.x;
                if (t >= num_tiles) t = -1;
                mbar_wait(qempty_bar(q), ((it >> 2) & 1) ^ 1);
                q_tile[q] = t;
                mbar_arrive(qfull_bar(q));                  // release: the tile index is visible to the consumers
                if (t < 0) break;
                int m0, n0;
                tile_coords(t, num_m, num_n, m0, n0);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1);
                    const uint32_t sa = base + stage * Cfg::stage_bytes;
                    mbar_expect_tx(full_bar(stage), Cfg::stage_bytes);
                    tma_load_2d(&tm_ahi, full_bar(stage), sa, kb * BK, m0);
                    tma_load_2d(&tm_alo, full_bar(stage), sa + Cfg::a_bytes, kb * BK, m0);
                    tma_load_2d(&tm_bhi, full_bar(stage), sa + 2 * Cfg::a_bytes, kb * BK, n0);
                    tma_load_2d(&tm_blo, full_bar(stage), sa + 2 * Cfg::a_bytes + Cfg::b_bytes, kb * BK, n0);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            // instruction descriptor: D=F32, A=B=TF32, both K-major, N=256, M=128, optional negate of A
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((negate ? 1u : 0u) << 13) |
                                   ((uint32_t)(TBN >> 3) << 17) | ((uint32_t)(TBM >> 4) << 24);
            int stage = 0; uint32_t phase = 0;
            for (int it = 0;; ++it) {
                const int q = it & 3;
                mbar_wait(qfull_bar(q), (it >> 2) & 1);
                const int t = q_tile[q];
                mbar_arrive(qempty_bar(q));
                if (t < 0) break;
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                mbar_wait(tempty_bar(as), aphase ^ 1);
                tc_fence_after();
                const uint32_t tacc = tmem_base + (uint32_t)(as * TBN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sa = base + stage * Cfg::stage_bytes;
                    const uint64_t d_ahi = make_kmajor_desc<Cfg::swizzle_bytes>(sa);
                    const uint64_t d_alo = make_kmajor_desc<Cfg::swizzle_bytes>(sa + Cfg::a_bytes);
                    const uint64_t d_bhi = make_kmajor_desc<Cfg::swizzle_bytes>(sa + 2 * Cfg::a_bytes);
                    const uint64_t d_blo = make_kmajor_desc<Cfg::swizzle_bytes>(sa + 2 * Cfg::a_bytes + Cfg::b_bytes);
#pragma unroll
                    for (int kk = 0; kk < BK / 8; ++kk) {
                        const uint64_t adv = (uint64_t)((kk * 32) >> 4);      // 8 tf32 = 32 B along K inside the swizzle row
                        umma_tf32(tacc, d_alo + adv, d_bhi + adv, idesc, (kb | kk) ? 1u : 0u);
                        umma_tf32(tacc, d_ahi + adv, d_blo + adv, idesc, 1u);
                        umma_tf32(tacc, d_ahi + adv, d_bhi + adv, idesc, 1u);
                    }
                    umma_commit(empty_bar(stage));          // frees the smem slot when these MMAs retire
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(tfull_bar(as));                 // accumulator ready for the epilogue
            }
        }
    } else {
        // ===================== epilogue (warps 2..5, 128 threads) =====================
        const int q = warp & 3;                             // TMEM lane quarter this warp may touch
        const int row = q * 32 + lane;
        const bool issuer = (threadIdx.x == 64);
        int cbuf = 0;
        for (int it = 0;; ++it) {
            const int q = it & 3;
            mbar_wait(qfull_bar(q), (it >> 2) & 1);
            const int t = q_tile[q];
            mbar_arrive(qempty_bar(q));
            if (t < 0) break;
            int m0, n0;
            tile_coords(t, num_m, num_n, m0, n0);
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            mbar_wait(tfull_bar(as), aphase);
            tc_fence_after();
            const int nchunks = min(TBN / TEPI_W, (n - n0 + TEPI_W - 1) / TEPI_W);
            for (int c = 0; c < nchunks; ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * TBN + c * TEPI_W), v);
                tmem_ld_wait();
                if (c == nchunks - 1) {                      // all of this accumulator has been read
                    tc_fence_before();
                    mbar_arrive(tempty_bar(as));
                }
                const uint32_t sbuf = base + Cfg::ring_bytes + cbuf * TSTAGE_C;
                if (issuer) tma_wait_read<1>();             // the store that last used this buffer has read it
                named_bar_sync(1, 128);
                const uint32_t rbase = sbuf + row * 128;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t dst = rbase + ((uint32_t)(j ^ (row & 7)) << 4);     // 128 B swizzle, matches tm_c
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};"
                                 ::"r"(dst), "r"(v[4 * j]), "r"(v[4 * j + 1]), "r"(v[4 * j + 2]), "r"(v[4 * j + 3]) : "memory");
                }
                fence_async_smem();
                named_bar_sync(1, 128);
                if (issuer) {
                    if (reduce) tma_reduce_add_2d(&tm_c, sbuf, n0 + c * TEPI_W, m0);
                    else        tma_store_2d(&tm_c, sbuf, n0 + c * TEPI_W, m0);
                    tma_commit();
                }
                cbuf ^= 1;
            }
        }
        if (issuer) tma_wait_all();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
    // the last CTA to leave rearms the scheduler slot for a later launch (sched[0] tile counter, sched[1] exits)
    if (dynamic && threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(sched + 1, 1) == (int)gridDim.x - 1) { sched[0] = 0; sched[1] = 0; __threadfence(); }
    }
}

// ---------------------------------------------------------------------------- the CTA-pair kernel
// Same pipeline with tcgen05.mma.cta_group::2: the two CTAs of a cluster (one TPC) compute a 256 x 256 tile.
// Each CTA stages its own 128 rows of A and HALF of the B tile (128 of the 256 columns); the MMA, issued by
// one thread of the even CTA, reads both halves across the pair and accumulates rows 0..127 / 128..255 into the
// two CTAs' own TMEM.  Per CTA and k-slice that is 64 KB through TMA instead of 96 KB -- the L2 -> SM stream is
// what keeps the single-CTA kernel at ~80 % tensor-pipe utilisation -- and the smaller stage buys a third
// pipeline stage.  Hand-offs:
//   full[s]    (even CTA only)  expect_tx by the even producer for BOTH CTAs' bytes; both producers' TMA loads
//                               signal it (mbarrier address with the peer bit cleared)
//   empty[s]   (both)           tcgen05.commit ... multicast to both CTAs
//   tfull[a]   (both)           tcgen05.commit ... multicast: accumulator `a` complete
//   tempty[a]  (even CTA only)  256 arrivals: the 128 epilogue threads of each CTA (the odd CTA's arrive remotely)
// Tiles come from the same dynamic scheduler as in the single-CTA kernel; the even CTA's producer draws the tile
// and publishes it into both CTAs' queues (remote store + cluster-scope mbarrier arrive).  Each output element
// sees the same MMAs in the same order as in the single-CTA kernel, so results are bit-identical.
constexpr int TBM2 = 256;
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;            // shared::cluster address -> the even CTA of the pair

template <int BK, int STAGES>
struct TcCfg2 {
    static constexpr int a_bytes = TBM * BK * 4;           // this CTA's 128 rows of A
    static constexpr int b_bytes = (TBN / 2) * BK * 4;     // this CTA's half of B
    static constexpr int stage_bytes = 2 * a_bytes + 2 * b_bytes;
    static constexpr int ring_bytes = STAGES * stage_bytes;
    static constexpr int bar_off = ring_bytes + 2 * TSTAGE_C;
    static constexpr int smem_bytes = bar_off + 256 + 1024;
    static constexpr int swizzle_bytes = BK * 4;
};

__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap* map, uint32_t bar_even, uint32_t dst, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar_even), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void mbar_arrive_even(uint32_t bar)        // arrive on the even CTA's copy of `bar`
{
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & kPeerBitMask) : "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank)     // same smem offset in CTA `rank` of the cluster
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_cluster_u32(uint32_t cluster_addr, int v)
{
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(cluster_addr), "r"(v) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr)     // release at cluster scope: orders the remote store above
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cl(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cl(uint32_t bar, uint32_t parity)    // bounded, acquire at cluster scope
{
    if (mbar_try_wait_cl(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait_cl(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) {
            printf("tc_gemm_pair: mbarrier wait timed out (block %d thread %d bar 0x%x parity %u)\n", blockIdx.x, threadIdx.x, bar, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar)        // arrives on `bar` in both CTAs
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}

__device__ __forceinline__ void tile_coords2(int t, int num_m, int num_n, int& m0, int& n0)
{
    const int band_tiles = num_m * kBandN;
    const int band = t / band_tiles;
    const int in_band = t - band * band_tiles;
    const int wb = min(kBandN, num_n - band * kBandN);
    const int mi = in_band / wb;
    const int ni = band * kBandN + (in_band - mi * wb);
    m0 = mi * TBM2;
    n0 = ni * TBN;
}

template <int BK, int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTcThreads, 1)
tc_gemm_pair_kernel(const __grid_constant__ CUtensorMap tm_ahi, const __grid_constant__ CUtensorMap tm_alo,
                    const __grid_constant__ CUtensorMap tm_bhi, const __grid_constant__ CUtensorMap tm_blo,
                    const __grid_constant__ CUtensorMap tm_c,
                    int m, int n, int k, int negate, int reduce, int* __restrict__ sched, int dynamic, int max_tiles)
{
    using Cfg = TcCfg2<BK, STAGES>;
    extern __shared__ unsigned char smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;
    const uint32_t bar_base = base + Cfg::bar_off;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    // tile queue (as in the single-CTA kernel), one copy per CTA: the even CTA's producer draws the tile and
    // publishes it to both; qempty lives in the even CTA and collects every consumer of both CTAs
    auto qfull_bar = [&](int q) { return bar_base + 8u * (2 * STAGES + 5 + q); };
    auto qempty_bar = [&](int q) { return bar_base + 8u * (2 * STAGES + 9 + q); };
    const uint32_t q_tile_u32 = bar_base + 8u * (2 * STAGES + 13);
    unsigned char* gen_base = smem_dyn + (base - smem_u32(smem_dyn));
    volatile int* q_tile = reinterpret_cast<volatile int*>(gen_base + Cfg::bar_off + 8 * (2 * STAGES + 13));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();                // 0: even CTA (issues the MMAs), 1: its peer
    pdl_release();
    const int pair = (int)blockIdx.x >> 1, num_pairs = (int)gridDim.x >> 1;
    const int num_n = (n + TBN - 1) / TBN;
    const int num_m = (m + TBM2 - 1) / TBM2;
    const int num_tiles = num_m * num_n;
    const int num_kb = k / BK;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 256); }
        // consumers of a queue slot: even CTA = MMA issuer + 128 epilogue threads, odd CTA = producer + 128 epilogue threads
        for (int q = 0; q < 4; ++q) { mbar_init(qfull_bar(q), 1); mbar_init(qempty_bar(q), 2 * (1 + 128)); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {                                        // both CTAs, same warp: the pair allocates together
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync_all();                                     // barriers of both CTAs initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(gen_base + Cfg::bar_off + 8 * (2 * STAGES + 4));
    pdl_wait();

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (elect_one()) {
            int stage = 0; uint32_t phase = 0;
            for (int it = 0;; ++it) {
                const int q = it & 3;
                int t;
                if (rank == 0) {
                    // same policy as the single-CTA kernel: dynamic == 1 first tile static, the rest from the counter;
                    // dynamic == 2 (bounded CTAs) every tile from the counter; 0 static round-robin over the pairs
                    if (max_tiles > 0 && it >= max_tiles) t = -1;
                    else if (dynamic == 2) t = atomicAdd(sched, 1);
                    else if (it == 0 || !dynamic) t = pair + it * num_pairs;
                    else t = atomicAdd(sched, 1) + num_pairs;
                    if (t >= num_tiles) t = -1;
                    mbar_wait_cl(qempty_bar(q), ((it >> 2) & 1) ^ 1);
                    q_tile[q] = t;
                    st_cluster_u32(mapa_u32(q_tile_u32 + 4u * q, 1), t);
                    mbar_arrive(qfull_bar(q));
                    mbar_arrive_cluster(mapa_u32(qfull_bar(q), 1));
                } else {
                    mbar_wait_cl(qfull_bar(q), (it >> 2) & 1);
                    t = q_tile[q];
                    mbar_arrive_cluster(mapa_u32(qempty_bar(q), 0));
                }
                if (t < 0) break;
                int m0, n0;
                tile_coords2(t, num_m, num_n, m0, n0);
                const int am = m0 + (int)rank * TBM, bn = n0 + (int)rank * (TBN / 2);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1);
                    const uint32_t sa = base + stage * Cfg::stage_bytes;
                    const uint32_t fb = full_bar(stage) & kPeerBitMask;
                    if (rank == 0) mbar_expect_tx(full_bar(stage), 2 * Cfg::stage_bytes);
                    tma_load_2d_pair(&tm_ahi, fb, sa, kb * BK, am);
                    tma_load_2d_pair(&tm_alo, fb, sa + Cfg::a_bytes, kb * BK, am);
                    tma_load_2d_pair(&tm_bhi, fb, sa + 2 * Cfg::a_bytes, kb * BK, bn);
                    tma_load_2d_pair(&tm_blo, fb, sa + 2 * Cfg::a_bytes + Cfg::b_bytes, kb * BK, bn);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (even CTA only) =====================
        if (rank == 0 && elect_one()) {
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((negate ? 1u : 0u) << 13) |
                                   ((uint32_t)(TBN >> 3) << 17) | ((uint32_t)(TBM2 >> 4) << 24);
            int stage = 0; uint32_t phase = 0;
            for (int it = 0;; ++it) {
                const int q = it & 3;
                mbar_wait_cl(qfull_bar(q), (it >> 2) & 1);
                const int t = q_tile[q];
                mbar_arrive_cluster(mapa_u32(qempty_bar(q), 0));
                if (t < 0) break;
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                mbar_wait(tempty_bar(as), aphase ^ 1);
                tc_fence_after();
                const uint32_t tacc = tmem_base + (uint32_t)(as * TBN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sa = base + stage * Cfg::stage_bytes;
                    const uint64_t d_ahi = make_kmajor_desc<Cfg::swizzle_bytes>(sa);
                    const uint64_t d_alo = make_kmajor_desc<Cfg::swizzle_bytes>(sa + Cfg::a_bytes);
                    const uint64_t d_bhi = make_kmajor_desc<Cfg::swizzle_bytes>(sa + 2 * Cfg::a_bytes);
                    const uint64_t d_blo = make_kmajor_desc<Cfg::swizzle_bytes>(sa + 2 * Cfg::a_bytes + Cfg::b_bytes);
#pragma unroll
                    for (int kk = 0; kk < BK / 8; ++kk) {
                        const uint64_t adv = (uint64_t)((kk * 32) >> 4);
                        umma_tf32_pair(tacc, d_alo + adv, d_bhi + adv, idesc, (kb | kk) ? 1u : 0u);
                        umma_tf32_pair(tacc, d_ahi + adv, d_blo + adv, idesc, 1u);
                        umma_tf32_pair(tacc, d_ahi + adv, d_bhi + adv, idesc, 1u);
                    }
                    umma_commit_pair(empty_bar(stage));
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit_pair(tfull_bar(as));
            }
        }
    } else {
        // ===================== epilogue (warps 2..5 of both CTAs: each CTA drains its own 128 rows) =====================
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const bool issuer = (threadIdx.x == 64);
        int cbuf = 0;
        const uint32_t qe0 = mapa_u32(qempty_bar(0), 0);
        for (int it = 0;; ++it) {
            const int qs = it & 3;
            mbar_wait_cl(qfull_bar(qs), (it >> 2) & 1);
            const int t = q_tile[qs];
            mbar_arrive_cluster(qe0 + 8u * qs);
            if (t < 0) break;
            int m0, n0;
            tile_coords2(t, num_m, num_n, m0, n0);
            m0 += (int)rank * TBM;
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            mbar_wait(tfull_bar(as), aphase);
            tc_fence_after();
            const int nchunks = min(TBN / TEPI_W, (n - n0 + TEPI_W - 1) / TEPI_W);
            const bool live = m0 < m;                       // the odd CTA of the last tile row may hold no rows
            for (int c = 0; c < nchunks; ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * TBN + c * TEPI_W), v);
                tmem_ld_wait();
                if (c == nchunks - 1) {
                    tc_fence_before();
                    mbar_arrive_even(tempty_bar(as));
                }
                if (!live) continue;
                const uint32_t sbuf = base + Cfg::ring_bytes + cbuf * TSTAGE_C;
                if (issuer) tma_wait_read<1>();
                named_bar_sync(1, 128);
                const uint32_t rbase = sbuf + row * 128;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t dst = rbase + ((uint32_t)(j ^ (row & 7)) << 4);
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};"
                                 ::"r"(dst), "r"(v[4 * j]), "r"(v[4 * j + 1]), "r"(v[4 * j + 2]), "r"(v[4 * j + 3]) : "memory");
                }
                fence_async_smem();
                named_bar_sync(1, 128);
                if (issuer) {
                    if (reduce) tma_reduce_add_2d(&tm_c, sbuf, n0 + c * TEPI_W, m0);
                    else        tma_store_2d(&tm_c, sbuf, n0 + c * TEPI_W, m0);
                    tma_commit();
                }
                cbuf ^= 1;
            }
        }
        if (issuer) tma_wait_all();
    }
    tc_fence_before();
    cluster_sync_all();                                     // both CTAs are done with both TMEMs and all barriers
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
    // the last CTA to leave rearms the scheduler slot for a later launch (sched[0] tile counter, sched[1] exits)
    if (dynamic && threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(sched + 1, 1) == (int)gridDim.x - 1) { sched[0] = 0; sched[1] = 0; __threadfence(); }
    }
}

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static int g_num_sms = 0;
// Scheduler slots: {tile counter, exit counter} pairs in a file-scope device array (zero at load, rearmed by
// each launch's last CTA).  Launches rotate through the slots so that kernels in flight on different
// streams never share one.  A static array rather than cudaMalloc: matmul_cuda_device_noalloc relies on it.
constexpr int kSchedSlots = 64;
static __device__ int g_sched_slots[2 * kSchedSlots];
static int* g_sched_addr[16] = {nullptr};
static std::atomic<unsigned> g_sched_next{0};   // (contexts of a single-process group launch from several host threads)
// Callers that replay their launches from a CUDA graph own their scheduler slots (so two graphs running
// on different streams can never share one); everyone else uses the library's ring.
static thread_local int* t_sched_override = nullptr;
static thread_local int t_max_tiles = 0;
static thread_local int t_reserve_sms = 0;   // SMs this thread's following launches leave free (bulk updates beside a latency-bound panel chain)
constexpr int g_dynamic = 1;         // tiles come from the dynamic scheduler (static round-robin was measured and dropped)
static int g_reserve_sms = 0;        // SMs left free for concurrent kernels (NCCL broadcast of the look-ahead panel)
static int g_pair = 1;               // GAUSS_CUDA_TC_PAIR=0 switches the CTA-pair kernel off
static int g_pair_min_tiles = 0;     // GAUSS_CUDA_TC_PAIR_MIN_TILES: pair tiles from which a launch uses it (0: two rounds)
static thread_local bool t_pair_ok = true;   // callers that need the dynamic tile scheduler (multi-GPU) veto it

bool tc_available()
{
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return false;
    return major == 10;
}

static bool g_env_read = false;

int tc_setup_device(bool reread_env)
{
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) return GAUSS_CUDA_CUDA_ERR;
        g_encode = (EncodeTiledFn)fn;
    }
    if (reread_env || !g_env_read) {       // context creation re-reads the switches (tests flip them); other callers leave them
        const char* pr = getenv("GAUSS_CUDA_TC_PAIR");
        g_pair = pr ? (atoi(pr) ? 1 : 0) : 1;
        const char* pm = getenv("GAUSS_CUDA_TC_PAIR_MIN_TILES");
        g_pair_min_tiles = pm ? atoi(pm) : 0;
        const char* rs = getenv("GAUSS_CUDA_TC_RESERVE_SMS");
        g_reserve_sms = rs ? (atoi(rs) < 0 ? 0 : (atoi(rs) > 64 ? 64 : atoi(rs))) : 0;
        g_env_read = true;
    }
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (dev >= 0 && dev < 16 && !g_sched_addr[dev]) {
        int* p = nullptr;
        if (cudaGetSymbolAddress((void**)&p, g_sched_slots) != cudaSuccess) return GAUSS_CUDA_CUDA_ERR;
        g_sched_addr[dev] = p;
    }
    cudaError_t e = cudaFuncSetAttribute(tc_gemm_kernel<32, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<32, 2>::smem_bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(tc_gemm_pair_kernel<32, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg2<32, 3>::smem_bytes);
    return e == cudaSuccess ? GAUSS_CUDA_OK : GAUSS_CUDA_CUDA_ERR;
}

static int make_map(CUtensorMap* map, const float* ptr, uint64_t inner, uint64_t outer, uint64_t ld_elems,
                    uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle swz)
{
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {ld_elems * sizeof(float)};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptr, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

template <int BK, int STAGES>
static int launch_variant(const float* Ahi, const float* Alo, int lda, const float* Bhi, const float* Blo, int ldb,
                          float* C, int ldc, int m, int n, int k, int sign, int beta, cudaStream_t s)
{
    using Cfg = TcCfg<BK, STAGES>;
    const CUtensorMapSwizzle swz = BK == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    CUtensorMap ta, tal, tb, tbl, tc;
    int rc = 0;
    rc |= make_map(&ta, Ahi, k, m, lda, BK, TBM, swz);
    rc |= make_map(&tal, Alo, k, m, lda, BK, TBM, swz);
    rc |= make_map(&tb, Bhi, k, n, ldb, BK, TBN, swz);
    rc |= make_map(&tbl, Blo, k, n, ldb, BK, TBN, swz);
    rc |= make_map(&tc, C, n, m, ldc, TEPI_W, TBM, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return (int)cudaErrorInvalidValue;
    const int tiles = ((m + TBM - 1) / TBM) * ((n + TBN - 1) / TBN);
    const int avail = g_num_sms - std::max(g_reserve_sms, t_reserve_sms);
    int grid = tiles < avail ? tiles : avail;
    int dyn = (g_dynamic && tiles > grid) ? 1 : 0;
    int max_tiles = 0;
    if (t_max_tiles > 0 && g_dynamic && tiles > avail * t_max_tiles / 2) {     // bounded CTAs: only worth it for big launches
        max_tiles = t_max_tiles;
        grid = (tiles + max_tiles - 1) / max_tiles;
        dyn = 2;
    }
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 16 || !g_sched_addr[dev]) return (int)cudaErrorInvalidValue;
    int* sched = (t_sched_override ? t_sched_override : g_sched_addr[dev]) + 2 * (g_sched_next.fetch_add(1, std::memory_order_relaxed) % kSchedSlots);
    launch_k(tc_gemm_kernel<BK, STAGES>, dim3(grid), dim3(kTcThreads), (size_t)Cfg::smem_bytes, s, ta, tal, tb, tbl, tc, m, n, k, sign < 0 ? 1 : 0, beta ? 1 : 0, sched, dyn, max_tiles);
    ++g_launch_count;
    return (int)cudaGetLastError();
}

// CTA pairs: 256 x 256 tiles over floor(SMs / 2) pairs, same scheduling policy as launch_variant.  Used for launches
// with at least two rounds of pair tiles.
static int launch_pair(const float* Ahi, const float* Alo, int lda, const float* Bhi, const float* Blo, int ldb,
                       float* C, int ldc, int m, int n, int k, int sign, int beta, cudaStream_t s)
{
    constexpr int BK = 32, STAGES = 3;
    using Cfg = TcCfg2<BK, STAGES>;
    CUtensorMap ta, tal, tb, tbl, tc;
    int rc = 0;
    rc |= make_map(&ta, Ahi, k, m, lda, BK, TBM, CU_TENSOR_MAP_SWIZZLE_128B);
    rc |= make_map(&tal, Alo, k, m, lda, BK, TBM, CU_TENSOR_MAP_SWIZZLE_128B);
    rc |= make_map(&tb, Bhi, k, n, ldb, BK, TBN / 2, CU_TENSOR_MAP_SWIZZLE_128B);
    rc |= make_map(&tbl, Blo, k, n, ldb, BK, TBN / 2, CU_TENSOR_MAP_SWIZZLE_128B);
    rc |= make_map(&tc, C, n, m, ldc, TEPI_W, TBM, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return (int)cudaErrorInvalidValue;
    const int tiles = ((m + TBM2 - 1) / TBM2) * ((n + TBN - 1) / TBN);
    const int avail = (g_num_sms - std::max(g_reserve_sms, t_reserve_sms)) / 2;
    int pairs = std::min(tiles, avail);
    int dyn = (g_dynamic && tiles > pairs) ? 1 : 0;
    int max_tiles = 0;
    if (t_max_tiles > 0 && g_dynamic && tiles > avail * t_max_tiles / 2) {     // bounded CTAs, as in launch_variant
        max_tiles = t_max_tiles;
        pairs = (tiles + max_tiles - 1) / max_tiles;
        dyn = 2;
    }
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 16 || !g_sched_addr[dev]) return (int)cudaErrorInvalidValue;
    int* sched = (t_sched_override ? t_sched_override : g_sched_addr[dev]) + 2 * (g_sched_next.fetch_add(1, std::memory_order_relaxed) % kSchedSlots);
    launch_k(tc_gemm_pair_kernel<BK, STAGES>, dim3(2 * pairs), dim3(kTcThreads), (size_t)Cfg::smem_bytes, s, ta, tal, tb, tbl, tc, m, n, k, sign < 0 ? 1 : 0, beta ? 1 : 0,
             sched, dyn, max_tiles);
    ++g_launch_count;
    return (int)cudaGetLastError();
}

void tc_set_pair_ok(bool ok) { t_pair_ok = ok; }
void tc_set_sched_slots(int* slots) { t_sched_override = slots; }
void tc_set_max_tiles(int max_tiles) { t_max_tiles = max_tiles; }
void tc_set_reserve_sms(int n) { t_reserve_sms = n < 0 ? 0 : (n > 100 ? 100 : n); }

int tc_gemm_nt(const float* Ahi, const float* Alo, int lda, const float* Bhi, const float* Blo, int ldb,
               float* C, int ldc, int m, int n, int k, int sign, int beta, cudaStream_t s)
{
    if (m <= 0 || n <= 0 || k <= 0) return 0;
    if (!g_encode || (k % 32) != 0) return (int)cudaErrorInvalidValue;
    if (g_pair && t_pair_ok &&
        ((m + TBM2 - 1) / TBM2) * ((n + TBN - 1) / TBN) >= (g_pair_min_tiles > 0 ? g_pair_min_tiles : 2 * (g_num_sms / 2)))
        return launch_pair(Ahi, Alo, lda, Bhi, Blo, ldb, C, ldc, m, n, k, sign, beta, s);
    return launch_variant<32, 2>(Ahi, Alo, lda, Bhi, Blo, ldb, C, ldc, m, n, k, sign, beta, s);
}

} // namespace gb200
