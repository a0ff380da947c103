// libgauss_cuda.so -- the C ABI (include/gauss_cuda.h) and the host-side schedule of the blocked
// no-pivot LU.  The arithmetic lives in tc_gemm.cu (tcgen05 tile engine: trailing update), lu_panel.cu (diagonal
// factor, panel solves), lu_simt.cu (FFMA variant, back substitution, helpers) and lu_pivot.cu (pivoting panel).
//
// Algorithm (two-level right-looking Crout LU on the augmented matrix [A | b], FP32):
//   outer block K = rows/cols [k0, k0+kb), kb <= NB (NB a multiple of 128), owner rank = K mod world
//   P1 (owner)  for each 128-wide inner block j of the row panel:
//                 diag factor -> U-row solve over all columns to the right (RHS column included)
//                 -> L-col solve + rank-128 update of the panel rows below
//   bcast       packed row panel [kb x (ld-k0)] to every rank (world > 1)         [gauss_mpi :137]
//   P2 (all)    for each inner block j: L-col solve of the local rows below the panel,
//                 rank-128 update of their columns inside [k0, k0+kb)
//   P3 (all)    local rows below, columns right of the panel:  A22 -= L21 * U12   (K = kb)
// which reproduces computeGauss (gauss_internal_input.c:118-138) up to rounding order; back
// substitution (solveGauss :144-159) runs block-wise from the last outer block up.
// Schedules: factor_single (one GPU, captured into a CUDA graph; P2 and the wide part of P1 on side streams),
// factor_streamed (one GPU, host data: upload overlapped, up-looking across row chunks), factor_cp (multi-GPU, 3xTF32:
// the critical-path schedule -- diagonal region in an extended scratch that yields inv(L11) / inv(U11), panel solves as
// single GEMMs, the next panel's needs pushed into the next owner's memory, near / far U12 broadcasts; lu_cp.cu),
// factor_lookahead (multi-GPU, FFMA mode: the round-1 pipeline), factor_pivot (external-input variant),
// gauss_cuda_solve_refined (double-precision answers).  Multi-GPU contexts are either one rank per process
// (gauss_cuda_init_dist) or a group of rank contexts inside one process, driven by library threads (gauss_cuda_init).
#include "gauss_cuda.h"
#include "gauss_internal.cuh"

#include <dlfcn.h>
#include <nccl.h>       // types only; the library is loaded lazily with dlopen (single-GPU users never need it)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <algorithm>
#include <string>
#include <chrono>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

using namespace gb200;

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int cuda_fail(cudaError_t e, const char* what, int line)
{
    snprintf(g_err, sizeof g_err, "%s (gauss_cuda.cu:%d): %s", what, line, cudaGetErrorString(e));
    return GAUSS_CUDA_CUDA_ERR;
}
#define CK(call)                                                            \
    do {                                                                    \
        cudaError_t e_ = (call);                                            \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call, __LINE__);       \
    } while (0)

extern "C" const char* gauss_cuda_last_error(void) { return g_err; }

extern "C" const char* gauss_cuda_strerror(int st)
{
    switch (st) {
        case GAUSS_CUDA_OK:         return "ok";
        case GAUSS_CUDA_ZERO_PIVOT: return "The matrix is singular";   // the reference's message, gauss_internal_input.c:87
        case GAUSS_CUDA_BAD_ARG:    return "bad argument";
        case GAUSS_CUDA_CUDA_ERR:   return "CUDA error";
        case GAUSS_CUDA_NCCL_ERR:   return "NCCL error";
        case GAUSS_CUDA_NOT_READY:  return "no system resident on the device (call gauss_cuda_upload first)";
        default:                    return "unknown status";
    }
}

extern "C" int gauss_cuda_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// ------------------------------------------------------------------------------------------
// NCCL, loaded on demand
// ------------------------------------------------------------------------------------------
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;

static int nccl_load()
{
    if (g_nccl.handle) return GAUSS_CUDA_OK;
    const char* names[] = {"libnccl.so.2", "/usr/lib/x86_64-linux-gnu/libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* nm : names) { h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
    if (!h) { snprintf(g_err, sizeof g_err, "cannot load libnccl.so.2: %s", dlerror()); return GAUSS_CUDA_NCCL_ERR; }
    g_nccl.GetUniqueId    = (decltype(g_nccl.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank   = (decltype(g_nccl.CommInitRank))dlsym(h, "ncclCommInitRank");
    g_nccl.Broadcast      = (decltype(g_nccl.Broadcast))dlsym(h, "ncclBroadcast");
    g_nccl.AllGather      = (decltype(g_nccl.AllGather))dlsym(h, "ncclAllGather");
    g_nccl.CommDestroy    = (decltype(g_nccl.CommDestroy))dlsym(h, "ncclCommDestroy");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.Broadcast || !g_nccl.CommDestroy) {
        snprintf(g_err, sizeof g_err, "libnccl.so.2 lacks a required symbol");
        return GAUSS_CUDA_NCCL_ERR;
    }
    g_nccl.handle = h;
    return GAUSS_CUDA_OK;
}
static int nccl_fail(ncclResult_t r, const char* what, int line)
{
    snprintf(g_err, sizeof g_err, "%s (gauss_cuda.cu:%d): %s", what, line,
             g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "nccl failure");
    return GAUSS_CUDA_NCCL_ERR;
}
#define NK(call)                                                            \
    do {                                                                    \
        ncclResult_t r_ = (call);                                           \
        if (r_ != ncclSuccess) return nccl_fail(r_, #call, __LINE__);       \
    } while (0)

extern "C" int gauss_cuda_unique_id(void* id128)
{
    if (!id128) return GAUSS_CUDA_BAD_ARG;
    int rc = nccl_load();
    if (rc) return rc;
    ncclUniqueId id;
    NK(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return GAUSS_CUDA_OK;
}

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
struct gauss_cuda_ctx {
    int n = 0, np = 0, ld = 0, NB = 0, mode = 0;
    int rank = 0, world = 1, device = 0;
    int nblk = 0;            // outer blocks
    int rows_local = 0;      // rows of W held by this rank (multiple of 128)
    float* W = nullptr;      // working augmented matrix  [rows_local x ld]
    float* W0 = nullptr;     // resident pristine copy (gauss_cuda_upload), same shape
    float* pbuf[2] = {nullptr, nullptr};   // packed row panels [NB x ld], double buffered   (world > 1)
    float* la_hi[2] = {nullptr, nullptr};       // [rows_local x NB]  split L operand (3xTF32 mode);
    float* la_lo[2] = {nullptr, nullptr};       //                    second copy for the single-GPU look-ahead
    cudaStream_t s_panel = nullptr;             // higher-priority stream of the single-GPU look-ahead
    cudaStream_t s_side = nullptr;              // side stream of the single-GPU schedule: P2(j) beside the rest of P1
    cudaStream_t s_wide = nullptr;              // single GPU: the part of P1 right of the panel, beside its in-panel chain
    cudaEvent_t ev_wide = nullptr;
    bool wide_split = true;                     // GAUSS_CUDA_WIDE_SPLIT=0 keeps P1 on one stream
    cudaEvent_t ev_inner[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}, ev_side = nullptr;
    cudaEvent_t ev_fork[2] = {nullptr, nullptr}, ev_join[2] = {nullptr, nullptr};
    const bool overlap = false;                 // (a single-GPU outer look-ahead was measured in round 1 and dropped: no gain)
    int overlap_max_tiles = 3;
    bool dist_pair = true;                      // GAUSS_CUDA_DIST_PAIR=0: multi-GPU updates stay on the single-CTA kernel
    float* ub_hi[2] = {nullptr, nullptr};       // [ld x NB]          split U operand, transposed;
    float* ub_lo[2] = {nullptr, nullptr};       //                    second copy only when world > 1 (look-ahead)
    float* xvec = nullptr;   // [np] solution
    float* hvec = nullptr;   // [np] staging for RHS upload / B_out
    int* status = nullptr;   // device: 0 or (row+1) of the first bad pivot
    int* trsv_ctrl = nullptr;   // device: ticket + per-128-block flags of the fused back substitution
    int* sched = nullptr;       // device: this context's tile-scheduler slots for the TC GEMM (2 x 64 ints)
    // the whole elimination + back substitution of a single-GPU context is a fixed launch sequence:
    // it is captured once into a CUDA graph and replayed, which takes the host out of the loop
    // (at N = 2048 the ~90 launches otherwise cost more host time than the GPU needs to run them)
    cudaGraphExec_t gexec = nullptr;
    bool capturing = false, graph_failed = false, use_graph = true;
    bool last_graph = false;            // the last run_core replayed the captured graph (reported in gauss_cuda_stats)
    long long graph_launches = 0;
    int graph_upd_count = 0;
    bool pivot = false;         // partial pivoting (external-input variant)
    float* pt = nullptr;        // packed panel scratch [128 x np]
    int* ipiv = nullptr;        // device: pivot row of every step
    // upload overlapped with the elimination (gauss_cuda_solve on a single GPU, see factor_streamed)
    bool stream_on = true;      // GAUSS_CUDA_STREAM=0 switches it off
    int stream_min_n = 16384;   // GAUSS_CUDA_STREAM_MIN_N: smaller systems upload first and replay the graph
    int stream_chunks = 0;      // GAUSS_CUDA_STREAM_CHUNKS: > 0 = that many equal row chunks; 0 = graded (small first)
    cudaStream_t s_copy = nullptr;
    std::vector<cudaEvent_t> ev_chunk;
    cudaEvent_t ev_gate = nullptr, ev_h2d_done = nullptr;
    float* ub_all_hi = nullptr; // [nblk][ld x NB] split U operand of every panel (3xTF32 mode)
    float* ub_all_lo = nullptr;
    bool last_streamed = false;
    // iterative refinement (gauss_cuda_solve_refined), allocated on first use
    double* A64 = nullptr;      // [n x n] the system in double
    double* v64 = nullptr;      // b | x | r | 2 norms   (3n + 2 doubles)
    float* rf32 = nullptr;      // r | y | d             (3 np floats)
    cudaStream_t s_main = nullptr, s_comm = nullptr;
    cudaEvent_t ev_ready[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr}, ev_free[2] = {nullptr, nullptr};
    cudaEvent_t ev_diag[2] = {nullptr, nullptr}, ev_cols[2] = {nullptr, nullptr}, ev_p2[2] = {nullptr, nullptr};
    cudaEvent_t ev_rdiag[2] = {nullptr, nullptr};   // owner: the kb x kb diagonal region of the panel is final (in-panel chain done)
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool trace = false;                 // GAUSS_CUDA_TRACE=1: per-phase timestamps of the multi-GPU main stream
    std::vector<cudaEvent_t> tr_ev;     // 6 per outer block
    std::vector<cudaEvent_t> upd_ev;    // 2 per outer block: brackets of the trailing-update launches
    std::vector<double> upd_flops;
    int upd_count = 0;
    ncclComm_t comm = nullptr;
    // critical-path schedule (factor_cp): the diagonal region is factored in an extended scratch that also yields
    // inv(L11) and inv(U11); both panel solves are GEMMs; the small "crit" package travels ahead of the bulk U12
    bool cp = false;
    int cp_reserve = 16;                        // SMs the bulk update leaves to the panel chain (GAUSS_CUDA_CP_RESERVE)
    // peer-memory push of the crit package to the next owner (multi-GPU, one process per GPU: CUDA IPC mapping)
    bool p2p = false;
    float* pbox = nullptr;                      // this rank's receive box: [1024 B of flag words | pcrit[0] | pcrit[1]], pcrit = 4 x NB x NB floats
    float* peer_pbox = nullptr;                 // the next rank's box, mapped into this process
    float* peer_box[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // every other rank's box (world <= 8)
    float* iu_all = nullptr;                    // [owned blocks][NB x NB] inv(U11) of this rank's outer blocks (back substitution)
    unsigned bs_count = 0;                      // back substitutions run by this context
    unsigned* push_counters = nullptr;          // device: completion counters of the two pushing kernels
    unsigned solve_seq = 0;                     // factorisations run by this context (every rank counts alike): flag values are seq * nblk + K + 1
    bool memops = false;                        // cuStreamWaitValue32 available (else a spinning one-warp kernel waits)
    int cp_near = 8;                            // outer blocks after the crit block whose U12^T travels (and is applied) ahead of the rest (GAUSS_CUDA_CP_NEAR)
    cudaEvent_t ev_near[2] = {nullptr, nullptr}, ev_nearlocal[2] = {nullptr, nullptr}, ev_rownear[2] = {nullptr, nullptr};
    cudaEvent_t ev_nearsplit[2] = {nullptr, nullptr};
    bool pdl = true;                            // GAUSS_CUDA_PDL=0: the chain's kernels are launched without programmatic dependent launch
    double host_issue_ms = 0.0;                 // host time spent issuing the last factorisation (trace)
    bool cp_nobulk = false;                     // GAUSS_CUDA_CP_NOBULK=1: timing diagnostic only (WRONG results): the bulk update is skipped
    float* ext = nullptr;                       // [2NB x 2NB] extended scratch E
    float* ext_la[2] = {nullptr, nullptr};      // hi/lo [2NB x NB]  split L of the chain inside E
    float* ext_ub[2] = {nullptr, nullptr};      // hi/lo [2NB x NB]  split U of the chain inside E (transposed)
    float* il[2] = {nullptr, nullptr};          // hi/lo [NB x NB]   inv(L11), row-major
    float* crit[2] = {nullptr, nullptr};        // per parity: iut_hi | iut_lo | uc_hi | uc_lo, NB x NB each (what the next panel needs first)
    float* as_[2] = {nullptr, nullptr};         // hi/lo [rows_local x NB]  split of A21 before the solve (A operand of L21 = A21 inv(U11))
    float* a12t[2] = {nullptr, nullptr};        // hi/lo [ld x NB]   split of A12 transposed (operand of both U12 GEMMs)
    float* ut[2] = {nullptr, nullptr};          // per parity [ld x NB]  raw U12 transposed: the bulk broadcast payload
    cudaEvent_t ev_crit[2] = {nullptr, nullptr}, ev_critlocal[2] = {nullptr, nullptr}, ev_critfree[2] = {nullptr, nullptr};
    cudaEvent_t ev_ut[2] = {nullptr, nullptr}, ev_utlocal[2] = {nullptr, nullptr}, ev_utfree[2] = {nullptr, nullptr};
    cudaEvent_t ev_la[2] = {nullptr, nullptr}, ev_colstrip[2] = {nullptr, nullptr}, ev_rowrest[2] = {nullptr, nullptr};
    cudaEvent_t ev_inv[2] = {nullptr, nullptr}, ev_p4a = nullptr, ev_wjoin = nullptr;
    bool resident = false, factored = false;
    double bcast_bytes = 0.0;
    gauss_cuda_stats last{};
    // single-process group (gauss_cuda_init with ngpus > 1): this object is only a handle; members[r] is the context of
    // rank r on device r, each driven by its own host thread during a call
    std::vector<gauss_cuda_ctx*> members;
    bool in_group = false;                      // a member: peers are reached by direct peer access, not CUDA IPC
    std::vector<float> group_tmp;               // host scratch for the X / B_out of ranks > 0
};

// Runs `fn(member, rank)` for every member of a group, one host thread per rank (rank 0 on the caller's), and returns
// the first failing status; the failing thread's message becomes the caller's gauss_cuda_last_error().
static int group_call(gauss_cuda_ctx* g, const std::function<int(gauss_cuda_ctx*, int)>& fn)
{
    const int n = (int)g->members.size();
    std::vector<int> rc((size_t)n, GAUSS_CUDA_OK);
    std::vector<std::string> msg((size_t)n);
    std::vector<std::thread> th;
    for (int r = 1; r < n; ++r)
        th.emplace_back([&, r] { rc[r] = fn(g->members[r], r); if (rc[r]) msg[r] = g_err; });
    rc[0] = fn(g->members[0], 0);
    for (auto& t : th) t.join();
    for (int r = 0; r < n; ++r)
        if (rc[r]) { if (r) snprintf(g_err, sizeof g_err, "rank %d: %s", r, msg[r].c_str()); return rc[r]; }
    return GAUSS_CUDA_OK;
}

static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// rows of outer block R: [R*NB, min((R+1)*NB, np))
static inline int blk_rows(const gauss_cuda_ctx* c, int R) { return std::min(c->NB, c->np - R * c->NB); }
// local row offset of outer block R on its owner
static inline int blk_local_row(const gauss_cuda_ctx* c, int R) { return (R / c->world) * c->NB; }
// first local block index whose global block index is > K  (rows strictly below outer block K)
static inline int first_local_block_after(const gauss_cuda_ctx* c, int K) { return K >= c->rank ? (K - c->rank) / c->world + 1 : 0; }
// number of local blocks whose global block index is < K   (rows strictly above outer block K)
static inline int local_blocks_before(const gauss_cuda_ctx* c, int K) { return K > c->rank ? (K - c->rank - 1) / c->world + 1 : 0; }


// Stream memory operations (cuStreamWaitValue32), fetched through the runtime like the tensor-map encoder
typedef int (*StreamWaitValue32Fn)(cudaStream_t, unsigned long long, unsigned, unsigned);
static StreamWaitValue32Fn g_wait_value32 = nullptr;

// Maps the NEXT rank's receive box into this process (CUDA IPC) so the kernels that produce the crit package can store
// it there directly.  Collective over the context's communicator; when any rank cannot map its neighbour (no peer
// access, GAUSS_CUDA_P2P=0) every rank falls back to the broadcast alone.
static int p2p_setup(gauss_cuda_ctx* c)
{
    const size_t nb2 = (size_t)c->NB * c->NB;
    const size_t box_bytes = 1024 + 8 * nb2 * sizeof(float) + (size_t)c->np * sizeof(float);   // flags | pcrit[2] | solution receive vector
    CK(cudaMalloc((void**)&c->pbox, box_bytes));
    CK(cudaMemset(c->pbox, 0, box_bytes));
    CK(cudaMalloc((void**)&c->push_counters, 256));
    CK(cudaMemset(c->push_counters, 0, 256));
    if (!g_nccl.AllGather) return GAUSS_CUDA_OK;                   // old NCCL without the symbol: broadcast only
    unsigned char* stage = nullptr;
    const size_t slot = 128;
    CK(cudaMalloc((void**)&stage, slot * (c->world + 1)));
    std::vector<unsigned char> host(slot * (c->world + 1), 0);
    cudaIpcMemHandle_t h;
    int ok = cudaIpcGetMemHandle(&h, c->pbox) == cudaSuccess ? 1 : 0;
    cudaGetLastError();
    memcpy(host.data(), &h, sizeof h);
    CK(cudaMemcpy(stage, host.data(), slot, cudaMemcpyHostToDevice));
    NK(g_nccl.AllGather(stage, stage + slot, slot, ncclChar, c->comm, c->s_main));
    CK(cudaStreamSynchronize(c->s_main));
    CK(cudaMemcpy(host.data(), stage, slot * (c->world + 1), cudaMemcpyDeviceToHost));
    const char* env = getenv("GAUSS_CUDA_P2P");
    if (env && atoi(env) == 0) ok = 0;
    if (c->world > 8) ok = 0;
    for (int r = 0; ok && r < c->world; ++r) {                     // every peer's box (the crit push uses the next rank's only)
        if (r == c->rank) continue;
        cudaIpcMemHandle_t hn;
        memcpy(&hn, host.data() + slot * (1 + r), sizeof hn);
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, hn, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess) c->peer_box[r] = (float*)p;
        else { ok = 0; cudaGetLastError(); }
    }
    if (ok) c->peer_pbox = c->peer_box[(c->rank + 1) % c->world];
    memset(host.data(), 0, slot);
    host[0] = (unsigned char)ok;
    CK(cudaMemcpy(stage, host.data(), slot, cudaMemcpyHostToDevice));
    NK(g_nccl.AllGather(stage, stage + slot, slot, ncclChar, c->comm, c->s_main));
    CK(cudaStreamSynchronize(c->s_main));
    CK(cudaMemcpy(host.data(), stage, slot * (c->world + 1), cudaMemcpyDeviceToHost));
    bool all = true;
    for (int r = 0; r < c->world; ++r) all = all && host[slot * (1 + r)] == 1;
    CK(cudaFree(stage));
    c->p2p = all;
    if (all) {
        int nown = 0;
        for (int R = c->rank; R < c->nblk; R += c->world) ++nown;
        CK(cudaMalloc((void**)&c->iu_all, (size_t)std::max(nown, 1) * nb2 * sizeof(float)));
    }
    if (!g_wait_value32) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &q) == cudaSuccess && fn) g_wait_value32 = (StreamWaitValue32Fn)fn;
        cudaGetLastError();
    }
    const char* mo = getenv("GAUSS_CUDA_MEMOPS");
    c->memops = g_wait_value32 != nullptr && !(mo && atoi(mo) == 0);
    return GAUSS_CUDA_OK;
}

static int ctx_create(gauss_cuda_ctx** out, int n, int rank, int world, int device, const void* nccl_id, int nb, int mode, bool in_group = false)
{
    if (!out) return GAUSS_CUDA_BAD_ARG;
    *out = nullptr;
    if (n < 1 || world < 1 || rank < 0 || rank >= world) return GAUSS_CUDA_BAD_ARG;
    if (mode != GAUSS_CUDA_MODE_3XTF32 && mode != GAUSS_CUDA_MODE_FFMA) return GAUSS_CUDA_BAD_ARG;
    if (nb < 0 || (nb % kNb) != 0) return GAUSS_CUDA_BAD_ARG;
    if (gauss_cuda_device_count() <= device) {
        snprintf(g_err, sizeof g_err, "no CUDA device %d (this library has no CPU fallback)", device);
        return GAUSS_CUDA_CUDA_ERR;
    }
    CK(cudaSetDevice(device));
    if (mode == GAUSS_CUDA_MODE_3XTF32 && !tc_available()) {
        snprintf(g_err, sizeof g_err, "GAUSS_CUDA_MODE_3XTF32 needs an sm_100 device (tcgen05/TMA)");
        return GAUSS_CUDA_CUDA_ERR;
    }
    gauss_cuda_ctx* c = new gauss_cuda_ctx();
    c->n = n; c->rank = rank; c->world = world; c->device = device; c->mode = mode; c->in_group = in_group;
    { const char* g = getenv("GAUSS_CUDA_GRAPH"); if (g && atoi(g) == 0) c->use_graph = false; }
    { const char* pd = getenv("GAUSS_CUDA_PDL"); if (pd) c->pdl = atoi(pd) != 0; }
    {   // tiles a CTA of the bulk update takes before it gives its SM up (round-1 multi-GPU look-ahead)
        const char* mt = getenv("GAUSS_CUDA_OVERLAP_TILES");
        if (mt && atoi(mt) > 0) c->overlap_max_tiles = atoi(mt);
    }
    { const char* e = getenv("GAUSS_CUDA_DIST_PAIR"); if (e && atoi(e) == 0) c->dist_pair = false; }
    { const char* e = getenv("GAUSS_CUDA_WIDE_SPLIT"); if (e && atoi(e) == 0) c->wide_split = false; }
    { const char* e = getenv("GAUSS_CUDA_STREAM"); if (e && atoi(e) == 0) c->stream_on = false; }
    { const char* e = getenv("GAUSS_CUDA_STREAM_MIN_N"); if (e) c->stream_min_n = atoi(e); }
    { const char* e = getenv("GAUSS_CUDA_STREAM_CHUNKS"); if (e && atoi(e) > 0) c->stream_chunks = atoi(e); }
    {   // schedule: the critical-path form is the multi-GPU default in 3xTF32 mode; GAUSS_CUDA_SCHED=cp selects it on one
        // GPU too, GAUSS_CUDA_SCHED=lookahead keeps the round-1 multi-GPU pipeline (the FFMA mode always uses that one)
        const char* e = getenv("GAUSS_CUDA_SCHED");
        const bool want = e ? (strcmp(e, "cp") == 0) : (world > 1);
        c->cp = want && mode == GAUSS_CUDA_MODE_3XTF32;
        const char* r = getenv("GAUSS_CUDA_CP_RESERVE");
        if (r) c->cp_reserve = std::max(0, std::min(64, atoi(r)));
        if (!r) c->cp_reserve = world >= 4 ? 32 : 16;
        const char* nr = getenv("GAUSS_CUDA_CP_NEAR");
        if (nr) c->cp_near = std::max(2, std::min(64, atoi(nr)));     // >= 2: see the column bookkeeping in factor_cp
        const char* nbk = getenv("GAUSS_CUDA_CP_NOBULK");
        c->cp_nobulk = nbk && atoi(nbk) != 0;
    }
    c->np = round_up(n, kNb);
    c->ld = c->np + kPadCols;
    if (nb == 0) {
        if (mode == GAUSS_CUDA_MODE_FFMA) nb = (world > 1) ? 256 : 128;
        // measured on B200 (3xTF32, one GPU): N=2048 nb 128/256/512 -> 1.07/1.12/1.13 ms; N=8192 256/512 -> 5.52/5.20 ms;
        // N=16384 256/512/1024 -> 20.8/18.6/18.6 ms; N=32768 256/512/768/1024 -> 127.9/113.3/111.1/109.7 ms
        else nb = (c->np >= 24576 && world == 1) ? 1024 : c->np >= 8192 ? 512 : (c->np >= 4096 ? 256 : 128);
    }
    c->NB = std::min(nb, c->np);
    c->nblk = (c->np + c->NB - 1) / c->NB;
    int rows = 0;
    for (int R = rank; R < c->nblk; R += world) rows += blk_rows(c, R);
    c->rows_local = rows;
    simt_setup_device();
    if (mode == GAUSS_CUDA_MODE_3XTF32) { int rc = tc_setup_device(); if (rc) { delete c; return rc; } }

    const size_t wbytes = (size_t)std::max(rows, 1) * c->ld * sizeof(float);
    cudaError_t e = cudaSuccess;
    auto A = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
    A((void**)&c->W, wbytes);
    A((void**)&c->xvec, (size_t)c->np * sizeof(float));
    A((void**)&c->hvec, (size_t)c->np * sizeof(float));
    A((void**)&c->status, sizeof(int));
    A((void**)&c->trsv_ctrl, (size_t)(c->np / kNb + 1) * sizeof(int));
    A((void**)&c->sched, 128 * sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(c->sched, 0, 128 * sizeof(int));
    if (world > 1 && !c->cp) for (int i = 0; i < 2; ++i) A((void**)&c->pbuf[i], (size_t)c->NB * c->ld * sizeof(float));
    if (c->cp) {
        const size_t nb2 = (size_t)c->NB * c->NB * sizeof(float);
        A((void**)&c->ext, 4 * nb2);
        for (int i = 0; i < 2; ++i) {
            A((void**)&c->ext_la[i], 2 * nb2); A((void**)&c->ext_ub[i], 2 * nb2);
            A((void**)&c->il[i], nb2);
            A((void**)&c->crit[i], 4 * nb2);
            A((void**)&c->as_[i], (size_t)std::max(rows, 1) * c->NB * sizeof(float));
            A((void**)&c->a12t[i], (size_t)c->ld * c->NB * sizeof(float));
            A((void**)&c->ut[i], (size_t)c->ld * c->NB * sizeof(float));
            if (e == cudaSuccess) e = cudaMemset(c->il[i], 0, nb2);
            if (e == cudaSuccess) e = cudaMemset(c->crit[i], 0, 4 * nb2);
            if (e == cudaSuccess) e = cudaMemset(c->a12t[i], 0, (size_t)c->ld * c->NB * sizeof(float));
            if (e == cudaSuccess) e = cudaMemset(c->ut[i], 0, (size_t)c->ld * c->NB * sizeof(float));
            if (e == cudaSuccess) e = cudaMemset(c->as_[i], 0, (size_t)std::max(rows, 1) * c->NB * sizeof(float));
        }
    }
    if (mode == GAUSS_CUDA_MODE_3XTF32) {
        const size_t la = (size_t)std::max(rows, 1) * c->NB * sizeof(float);
        const size_t ub = (size_t)c->ld * c->NB * sizeof(float);
        A((void**)&c->la_hi[0], la); A((void**)&c->la_lo[0], la);
        if (c->overlap || world > 1 || c->cp) { A((void**)&c->la_hi[1], la); A((void**)&c->la_lo[1], la); }
        for (int i = 0; i < ((world > 1 || c->overlap || c->cp) ? 2 : 1); ++i) {
            A((void**)&c->ub_hi[i], ub); A((void**)&c->ub_lo[i], ub);
            if (e == cudaSuccess) e = cudaMemset(c->ub_hi[i], 0, ub);
            if (e == cudaSuccess) e = cudaMemset(c->ub_lo[i], 0, ub);
        }
    }
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->s_main, cudaStreamNonBlocking);
    const bool dist_fork = (world > 1 || c->cp) && !(getenv("GAUSS_CUDA_DIST_FORK") && atoi(getenv("GAUSS_CUDA_DIST_FORK")) == 0);
    if ((world == 1 && !(getenv("GAUSS_CUDA_INNER_FORK") && atoi(getenv("GAUSS_CUDA_INNER_FORK")) == 0)) || dist_fork) {
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->s_side, cudaStreamNonBlocking);
        if (c->wide_split) {
            int lo_p = 0, hi_p = 0;
            cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p);
            // multi-GPU: the wide part of the panel factorisation belongs to the (priority) panel path
            if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&c->s_wide, cudaStreamNonBlocking, (world > 1 || c->cp) ? hi_p : lo_p);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_wide, cudaEventDisableTiming);
        }
        for (int i = 0; i < 8 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&c->ev_inner[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_side, cudaEventDisableTiming);
    }
    if (c->overlap || dist_fork) {
        int lo_p = 0, hi_p = 0;
        cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p);
        if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&c->s_panel, cudaStreamNonBlocking, hi_p);
        for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
            e = cudaEventCreateWithFlags(&c->ev_fork[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_join[i], cudaEventDisableTiming);
        }
    }
    if (c->cp) {
        cudaEvent_t* evs[] = {&c->ev_crit[0], &c->ev_crit[1], &c->ev_critlocal[0], &c->ev_critlocal[1], &c->ev_critfree[0], &c->ev_critfree[1],
                              &c->ev_ut[0], &c->ev_ut[1], &c->ev_utlocal[0], &c->ev_utlocal[1], &c->ev_utfree[0], &c->ev_utfree[1],
                              &c->ev_la[0], &c->ev_la[1], &c->ev_colstrip[0], &c->ev_colstrip[1], &c->ev_rowrest[0], &c->ev_rowrest[1],
                              &c->ev_inv[0], &c->ev_inv[1], &c->ev_p4a, &c->ev_wjoin,
                              &c->ev_near[0], &c->ev_near[1], &c->ev_nearlocal[0], &c->ev_nearlocal[1], &c->ev_rownear[0], &c->ev_rownear[1],
                              &c->ev_nearsplit[0], &c->ev_nearsplit[1]};
        for (cudaEvent_t* ev : evs) if (e == cudaSuccess) e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
    }
    if (world > 1 || c->cp) {
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->s_comm, cudaStreamNonBlocking);
        for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
            e = cudaEventCreateWithFlags(&c->ev_ready[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_diag[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_rdiag[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_cols[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_p2[i], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_free[i], cudaEventDisableTiming);
        }
    }
    for (int i = 0; i < 6 && e == cudaSuccess; ++i) e = cudaEventCreate(&c->ev[i]);
    c->trace = (world > 1 || c->cp) && getenv("GAUSS_CUDA_TRACE") && atoi(getenv("GAUSS_CUDA_TRACE")) != 0;
    if (c->trace) {
        c->tr_ev.assign(8 * (size_t)c->nblk, nullptr);
        for (size_t i = 0; i < c->tr_ev.size() && e == cudaSuccess; ++i) e = cudaEventCreate(&c->tr_ev[i]);
    }
    c->upd_ev.assign(10 * (size_t)c->nblk, nullptr);           // the schedules split the update of a step in up to five launches
    c->upd_flops.assign(5 * (size_t)c->nblk, 0.0);
    for (size_t i = 0; i < c->upd_ev.size() && e == cudaSuccess; ++i) e = cudaEventCreate(&c->upd_ev[i]);
    if (e != cudaSuccess) { int rc = cuda_fail(e, "context allocation", __LINE__); gauss_cuda_free(c); return rc; }

    if (world > 1) {
        int rc = nccl_load();
        if (rc) { gauss_cuda_free(c); return rc; }
        ncclUniqueId id;
        memcpy(&id, nccl_id, sizeof id);
        ncclResult_t r = g_nccl.CommInitRank(&c->comm, world, id, rank);
        if (r != ncclSuccess) { rc = nccl_fail(r, "ncclCommInitRank", __LINE__); gauss_cuda_free(c); return rc; }
        if (c->cp && !c->in_group && (rc = p2p_setup(c))) { gauss_cuda_free(c); return rc; }
    }
    *out = c;
    return GAUSS_CUDA_OK;
}

// Peer wiring of a single-process group: every member gets its receive box and direct pointers to the others' (peer
// access between the devices enabled both ways); any failure leaves the whole group on the broadcast-only path.
static int group_p2p_setup(gauss_cuda_ctx* g)
{
    const int n = (int)g->members.size();
    bool ok = n <= 8 && !(getenv("GAUSS_CUDA_P2P") && atoi(getenv("GAUSS_CUDA_P2P")) == 0);
    for (int r = 0; r < n && ok; ++r) ok = g->members[r]->cp;
    if (!ok) return GAUSS_CUDA_OK;
    for (int r = 0; r < n; ++r) {
        gauss_cuda_ctx* c = g->members[r];
        CK(cudaSetDevice(c->device));
        const size_t nb2 = (size_t)c->NB * c->NB;
        const size_t box_bytes = 1024 + 8 * nb2 * sizeof(float) + (size_t)c->np * sizeof(float);
        CK(cudaMalloc((void**)&c->pbox, box_bytes));
        CK(cudaMemset(c->pbox, 0, box_bytes));
        CK(cudaMalloc((void**)&c->push_counters, 256));
        CK(cudaMemset(c->push_counters, 0, 256));
        for (int q = 0; q < n && ok; ++q) {
            if (q == r) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, c->device, g->members[q]->device) != cudaSuccess || !can) { ok = false; break; }
            cudaError_t e = cudaDeviceEnablePeerAccess(g->members[q]->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = false;
            cudaGetLastError();
        }
    }
    if (!ok) return GAUSS_CUDA_OK;
    for (int r = 0; r < n; ++r) {
        gauss_cuda_ctx* c = g->members[r];
        CK(cudaSetDevice(c->device));
        for (int q = 0; q < n; ++q) if (q != r) c->peer_box[q] = g->members[q]->pbox;
        c->peer_pbox = c->peer_box[(r + 1) % n];
        int nown = 0;
        for (int R = c->rank; R < c->nblk; R += c->world) ++nown;
        CK(cudaMalloc((void**)&c->iu_all, (size_t)std::max(nown, 1) * c->NB * c->NB * sizeof(float)));
        if (!g_wait_value32) {
            void* fn = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &q) == cudaSuccess && fn) g_wait_value32 = (StreamWaitValue32Fn)fn;
            cudaGetLastError();
        }
        const char* mo = getenv("GAUSS_CUDA_MEMOPS");
        c->memops = g_wait_value32 != nullptr && !(mo && atoi(mo) == 0);
        c->p2p = true;
    }
    return GAUSS_CUDA_OK;
}

extern "C" int gauss_cuda_init(gauss_cuda_ctx** ctx, int n, int ngpus, int nb, int mode)
{
    if (ngpus > 1) {
        // one process, ngpus devices (0 .. ngpus-1): one rank context per device, each call runs them on their own host
        // threads -- the reference's own multi-worker form is one program too (mpirun ... gauss_internal_input, README.txt:23)
        if (!ctx) return GAUSS_CUDA_BAD_ARG;
        *ctx = nullptr;
        if (gauss_cuda_device_count() < ngpus) {
            snprintf(g_err, sizeof g_err, "gauss_cuda_init: %d GPUs asked for, %d visible (no CPU fallback)", ngpus, gauss_cuda_device_count());
            return GAUSS_CUDA_CUDA_ERR;
        }
        char id[128];
        int rc = gauss_cuda_unique_id(id);
        if (rc) return rc;
        gauss_cuda_ctx* g = new gauss_cuda_ctx();
        g->n = n; g->world = ngpus; g->mode = mode;
        g->members.assign((size_t)ngpus, nullptr);
        rc = group_call(g, [&](gauss_cuda_ctx*, int r) { return ctx_create(&g->members[r], n, r, ngpus, r, id, nb, mode, true); });
        if (rc == GAUSS_CUDA_OK) rc = group_p2p_setup(g);
        if (rc) { gauss_cuda_free(g); return rc; }
        g->group_tmp.assign(2 * (size_t)n * (size_t)ngpus, 0.f);
        *ctx = g;
        return GAUSS_CUDA_OK;
    }
    if (ngpus != 1) {
        snprintf(g_err, sizeof g_err, "gauss_cuda_init: ngpus must be >= 1 (got %d)", ngpus);
        return GAUSS_CUDA_BAD_ARG;
    }
    int dev = 0;
    if (gauss_cuda_device_count() > 0) cudaGetDevice(&dev);
    return ctx_create(ctx, n, 0, 1, dev, nullptr, nb, mode);
}

extern "C" int gauss_cuda_init_dist(gauss_cuda_ctx** ctx, int n, int rank, int world, int device,
                                    const void* nccl_id_128, int nb, int mode)
{
    if (world > 1 && !nccl_id_128) return GAUSS_CUDA_BAD_ARG;
    return ctx_create(ctx, n, rank, world, device, nccl_id_128, nb, mode);
}

extern "C" void gauss_cuda_free(gauss_cuda_ctx* c)
{
    if (!c) return;
    if (!c->members.empty()) {                                   // a group handle: NCCL teardown wants every rank, so in threads
        std::vector<std::thread> th;
        for (gauss_cuda_ctx* m : c->members) if (m) th.emplace_back([m] { gauss_cuda_free(m); });
        for (auto& t : th) t.join();
        delete c;
        return;
    }
    cudaSetDevice(c->device);
    if (c->s_main) cudaStreamSynchronize(c->s_main);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    if (c->s_comm) cudaStreamSynchronize(c->s_comm);
    if (c->s_panel) cudaStreamSynchronize(c->s_panel);
    if (c->s_side) cudaStreamSynchronize(c->s_side);
    if (c->s_wide) cudaStreamSynchronize(c->s_wide);
    float* bufs[] = {c->W, c->W0, c->pbuf[0], c->pbuf[1], c->la_hi[0], c->la_lo[0], c->la_hi[1], c->la_lo[1],
                     c->ub_hi[0], c->ub_lo[0], c->ub_hi[1], c->ub_lo[1], c->xvec, c->hvec};
    for (float* b : bufs) if (b) cudaFree(b);
    {
        float* cpb[] = {c->ext, c->ext_la[0], c->ext_la[1], c->ext_ub[0], c->ext_ub[1], c->il[0], c->il[1], c->crit[0], c->crit[1],
                        c->as_[0], c->as_[1], c->a12t[0], c->a12t[1], c->ut[0], c->ut[1]};
        for (float* b : cpb) if (b) cudaFree(b);
        cudaEvent_t evs[] = {c->ev_crit[0], c->ev_crit[1], c->ev_critlocal[0], c->ev_critlocal[1], c->ev_critfree[0], c->ev_critfree[1],
                             c->ev_ut[0], c->ev_ut[1], c->ev_utlocal[0], c->ev_utlocal[1], c->ev_utfree[0], c->ev_utfree[1],
                             c->ev_la[0], c->ev_la[1], c->ev_colstrip[0], c->ev_colstrip[1], c->ev_rowrest[0], c->ev_rowrest[1],
                             c->ev_inv[0], c->ev_inv[1], c->ev_p4a, c->ev_wjoin,
                             c->ev_near[0], c->ev_near[1], c->ev_nearlocal[0], c->ev_nearlocal[1], c->ev_rownear[0], c->ev_rownear[1],
                             c->ev_nearsplit[0], c->ev_nearsplit[1]};
        for (cudaEvent_t ev : evs) if (ev) cudaEventDestroy(ev);
    }
    if (!c->in_group) for (float* pb : c->peer_box) if (pb) cudaIpcCloseMemHandle(pb);
    if (c->iu_all) cudaFree(c->iu_all);
    if (c->pbox) cudaFree(c->pbox);
    if (c->push_counters) cudaFree(c->push_counters);
    if (c->status) cudaFree(c->status);
    if (c->gexec) cudaGraphExecDestroy(c->gexec);
    if (c->trsv_ctrl) cudaFree(c->trsv_ctrl);
    if (c->sched) cudaFree(c->sched);
    if (c->pt) cudaFree(c->pt);
    if (c->ipiv) cudaFree(c->ipiv);
    if (c->s_copy) { cudaStreamSynchronize(c->s_copy); cudaStreamDestroy(c->s_copy); }
    for (auto& e : c->ev_chunk) if (e) cudaEventDestroy(e);
    if (c->ev_gate) cudaEventDestroy(c->ev_gate);
    if (c->ev_h2d_done) cudaEventDestroy(c->ev_h2d_done);
    if (c->ub_all_hi) cudaFree(c->ub_all_hi);
    if (c->ub_all_lo) cudaFree(c->ub_all_lo);
    if (c->A64) cudaFree(c->A64);
    if (c->v64) cudaFree(c->v64);
    if (c->rf32) cudaFree(c->rf32);
    for (auto& e : c->ev) if (e) cudaEventDestroy(e);
    for (auto& e : c->upd_ev) if (e) cudaEventDestroy(e);
    for (auto& e : c->tr_ev) if (e) cudaEventDestroy(e);
    for (int i = 0; i < 2; ++i) {
        if (c->ev_ready[i]) cudaEventDestroy(c->ev_ready[i]);
        if (c->ev_done[i]) cudaEventDestroy(c->ev_done[i]);
        if (c->ev_diag[i]) cudaEventDestroy(c->ev_diag[i]);
        if (c->ev_rdiag[i]) cudaEventDestroy(c->ev_rdiag[i]);
        if (c->ev_cols[i]) cudaEventDestroy(c->ev_cols[i]);
        if (c->ev_p2[i]) cudaEventDestroy(c->ev_p2[i]);
        if (c->ev_free[i]) cudaEventDestroy(c->ev_free[i]);
    }
    if (c->s_comm) cudaStreamDestroy(c->s_comm);
    for (int i = 0; i < 2; ++i) {
        if (c->ev_fork[i]) cudaEventDestroy(c->ev_fork[i]);
        if (c->ev_join[i]) cudaEventDestroy(c->ev_join[i]);
    }
    if (c->s_panel) cudaStreamDestroy(c->s_panel);
    for (auto& ev : c->ev_inner) if (ev) cudaEventDestroy(ev);
    if (c->ev_side) cudaEventDestroy(c->ev_side);
    if (c->s_side) cudaStreamDestroy(c->s_side);
    if (c->ev_wide) cudaEventDestroy(c->ev_wide);
    if (c->s_wide) cudaStreamDestroy(c->s_wide);
    if (c->s_main) cudaStreamDestroy(c->s_main);
    delete c;
}

// ------------------------------------------------------------------------------------------
// host <-> device
// ------------------------------------------------------------------------------------------
// Copies this rank's rows of [A | b] into `dst` (same shape as W) and writes the identity padding.
// Outer blocks [R0, R1) of this rank; the RHS entries come from hvec (already on the device, same stream order).
static int upload_blocks(gauss_cuda_ctx* c, float* dst, const float* A, int R0, int R1, cudaStream_t s)
{
    const int n = c->n;
    for (int R = R0; R < R1; ++R) {
        if (R % c->world != c->rank) continue;
        const int g0 = R * c->NB, l0 = blk_local_row(c, R), rows = blk_rows(c, R);
        const int real = std::max(0, std::min(rows, n - g0));
        if (real > 0) {
            CK(cudaMemcpy2DAsync(dst + (size_t)l0 * c->ld, (size_t)c->ld * sizeof(float),
                                 A + (size_t)g0 * n, (size_t)n * sizeof(float),
                                 (size_t)n * sizeof(float), real, cudaMemcpyHostToDevice, s));
            launch_set_column(dst + (size_t)l0 * c->ld, c->ld, c->np, c->hvec + g0, real, s);
            // columns [n, ld) of real rows except the RHS column
            launch_pad_identity(dst, c->ld, l0, g0, real, n, n, s);
        }
        if (rows > real) launch_pad_identity(dst, c->ld, l0 + real, g0 + real, rows - real, n, 0, s);
    }
    CK(cudaGetLastError());
    return GAUSS_CUDA_OK;
}

static int upload_into(gauss_cuda_ctx* c, float* dst, const float* A, const float* B)
{
    CK(cudaMemcpyAsync(c->hvec, B, (size_t)c->n * sizeof(float), cudaMemcpyHostToDevice, c->s_main));
    return upload_blocks(c, dst, A, 0, c->nblk, c->s_main);
}

// ------------------------------------------------------------------------------------------
// elimination
// ------------------------------------------------------------------------------------------
static int gemm_update(gauss_cuda_ctx* c, cudaStream_t s, const float* A, int lda, const float* B, int ldb,
                       const float* Ahi, const float* Alo, const float* Bhi, const float* Blo,
                       float* C, int m, int ncols, int k, int max_tiles = 0, int reserve = 0)
{
    if (m <= 0 || ncols <= 0) return GAUSS_CUDA_OK;
    if (c->mode == GAUSS_CUDA_MODE_FFMA) {
        launch_sgemm_nn(A, lda, B, ldb, C, c->ld, m, ncols, k, -1.f, 1.f, s);
        return GAUSS_CUDA_OK;
    }
    tc_set_sched_slots(c->sched);
    tc_set_max_tiles(max_tiles);
    tc_set_pair_ok(c->world == 1 || c->dist_pair);
    tc_set_reserve_sms(reserve);
    int rc = tc_gemm_nt(Ahi, Alo, c->NB, Bhi, Blo, c->NB, C, c->ld, m, ncols, k, -1, 1, s);
    tc_set_reserve_sms(0);
    tc_set_max_tiles(0);
    tc_set_sched_slots(nullptr);
    if (rc) return cuda_fail((cudaError_t)rc, "tc_gemm_nt", __LINE__);
    return GAUSS_CUDA_OK;
}

// Event record that stays usable for cudaEventElapsedTime when the stream is being captured.
static cudaError_t rec_event(gauss_cuda_ctx* c, cudaEvent_t ev, cudaStream_t s)
{
    return c->capturing ? cudaEventRecordWithFlags(ev, s, cudaEventRecordExternal) : cudaEventRecord(ev, s);
}

// ---- the pieces of one outer step --------------------------------------------------------
struct PanelRef {            // where the factored row panel of the current outer block can be read
    const float* pd; int ldd;       // its diagonal region: element (panel row 0, column k0), kb columns
    const float* pu; int ldu;       // the U12 region:      element (panel row 0, column k0 + kb)
    float* ub_hi; float* ub_lo;     // split, transposed U operand of this step (3xTF32 mode)
};
struct StepBufs {            // the stream a piece runs on and the split-L buffers it reads / writes
    cudaStream_t s;
    float* la_hi; float* la_lo;
};

// P1: factor the row panel of outer block K in place on its owner.
static int panel_factor(gauss_cuda_ctx* c, int K, float* ub_hi, float* ub_lo, const StepBufs& sb,
                        cudaStream_t sw = nullptr, cudaEvent_t ev_in = nullptr)
{
    cudaStream_t s = sb.s;
    const int ld = c->ld, NB = c->NB, np = c->np;
    const bool tc = c->mode == GAUSS_CUDA_MODE_3XTF32;
    const int k0 = K * NB, kb = blk_rows(c, K), ninner = kb / kNb;
    const int lrow = blk_local_row(c, K);
    float* P = c->W + (size_t)lrow * ld;
    // split at the right edge of the panel (as in outer_step): the in-panel chain -- all that the kb x kb diagonal
    // region, and therefore the first broadcast and every rank's P2, depend on -- stays on `s`; the U-row solve and
    // the rank-128 update right of the panel follow on `sw`, one inner block behind.  ev_in fires when the diagonal
    // region is final; on return `s` has joined `sw` (the whole panel is done in stream order).
    const bool split = sw != nullptr && sw != s && ninner >= 2 && ninner <= 8;
    const int csplit = k0 + kb;
    for (int j = 0; j < ninner; ++j) {
        const int kk = k0 + j * kNb, r = j * kNb;
        float* D = P + (size_t)r * ld + kk;
        launch_diag_factor(D, ld, kk, c->status, s);
        const int mrem = kb - (r + kNb);
        // U-row solve and L-col solve of the panel rows below, one launch (both only need the factor)
        launch_urow_lcol(D, ld,
                         D + kNb, ld, split ? csplit - (kk + kNb) : np + 1 - (kk + kNb),
                         tc ? ub_hi + (size_t)(kk + kNb) * NB : nullptr, tc ? ub_lo + (size_t)(kk + kNb) * NB : nullptr, NB, r,
                         P + (size_t)(r + kNb) * ld + kk, ld, mrem,
                         tc ? sb.la_hi + (size_t)(lrow + r + kNb) * NB : nullptr, tc ? sb.la_lo + (size_t)(lrow + r + kNb) * NB : nullptr, NB, r, s);
        if (split) {
            CK(cudaEventRecord(c->ev_inner[j], s));
            CK(cudaStreamWaitEvent(sw, c->ev_inner[j], 0));
            launch_urow(D, ld, P + (size_t)r * ld + csplit, ld, np + 1 - csplit,
                        tc ? ub_hi + (size_t)csplit * NB : nullptr, tc ? ub_lo + (size_t)csplit * NB : nullptr, NB, r, sw);
            if (mrem > 0) {
                int rc = gemm_update(c, sw, P + (size_t)(r + kNb) * ld + kk, ld, P + (size_t)r * ld + csplit, ld,
                                     tc ? sb.la_hi + (size_t)(lrow + r + kNb) * NB + r : nullptr, tc ? sb.la_lo + (size_t)(lrow + r + kNb) * NB + r : nullptr,
                                     tc ? ub_hi + (size_t)csplit * NB + r : nullptr, tc ? ub_lo + (size_t)csplit * NB + r : nullptr,
                                     P + (size_t)(r + kNb) * ld + csplit, mrem, ld - csplit, kNb);
                if (rc) return rc;
            }
        }
        if (mrem > 0) {
            int rc = gemm_update(c, s, P + (size_t)(r + kNb) * ld + kk, ld, D + kNb, ld,
                                 tc ? sb.la_hi + (size_t)(lrow + r + kNb) * NB + r : nullptr, tc ? sb.la_lo + (size_t)(lrow + r + kNb) * NB + r : nullptr,
                                 tc ? ub_hi + (size_t)(kk + kNb) * NB + r : nullptr, tc ? ub_lo + (size_t)(kk + kNb) * NB + r : nullptr,
                                 P + (size_t)(r + kNb) * ld + kk + kNb, mrem, split ? csplit - (kk + kNb) : ld - (kk + kNb), kNb);
            if (rc) return rc;
        }
    }
    if (ev_in) CK(cudaEventRecord(ev_in, s));
    if (split) {
        CK(cudaEventRecord(c->ev_wide, sw));
        CK(cudaStreamWaitEvent(s, c->ev_wide, 0));
    }
    return GAUSS_CUDA_OK;
}

// Non-owners in 3xTF32 mode: rebuild the split, transposed U operand from the raw broadcast panel;
// `diag_part` selects the columns inside the panel (needed by P2) or right of it (needed by P3).
static void rebuild_split_u(gauss_cuda_ctx* c, int K, const PanelRef& pr, bool diag_part, cudaStream_t st)
{
    const int NB = c->NB, k0 = K * NB, kb = blk_rows(c, K), ninner = kb / kNb;
    for (int j = 0; j < ninner; ++j) {
        const int r = j * kNb;
        if (diag_part) {
            const int cfirst = r + kNb, ncol = kb - cfirst;        // panel-relative columns right of inner block j
            if (ncol > 0)
                launch_split_transpose(pr.pd + (size_t)r * pr.ldd + cfirst, pr.ldd,
                                       pr.ub_hi + (size_t)(k0 + cfirst) * NB + r, pr.ub_lo + (size_t)(k0 + cfirst) * NB + r,
                                       NB, kNb, ncol, kNb, st);
        } else {
            launch_split_transpose(pr.pu + (size_t)r * pr.ldu, pr.ldu,
                                   pr.ub_hi + (size_t)(k0 + kb) * NB + r, pr.ub_lo + (size_t)(k0 + kb) * NB + r,
                                   NB, kNb, c->ld - k0 - kb, kNb, st);
        }
    }
}

// P2: L-col solves and in-panel updates for the local rows [row_lo, row_hi)  (row_hi < 0: to the last local row).
static int col_panel(gauss_cuda_ctx* c, int K, const PanelRef& pr, int row_lo, const StepBufs& sb, int row_hi = -1)
{
    cudaStream_t s = sb.s;
    const int ld = c->ld, NB = c->NB;
    const bool tc = c->mode == GAUSS_CUDA_MODE_3XTF32;
    const int k0 = K * NB, kb = blk_rows(c, K), ninner = kb / kNb;
    const int m = (row_hi < 0 ? c->rows_local : row_hi) - row_lo;
    if (m <= 0) return GAUSS_CUDA_OK;
    float* Wb = c->W + (size_t)row_lo * ld;
    for (int j = 0; j < ninner; ++j) {
        const int kk = k0 + j * kNb, r = j * kNb;
        launch_lcol(pr.pd + (size_t)r * pr.ldd + r, pr.ldd, Wb + kk, ld, m,
                    tc ? sb.la_hi + (size_t)row_lo * NB : nullptr, tc ? sb.la_lo + (size_t)row_lo * NB : nullptr, NB, r, s);
        const int nin = kb - (r + kNb);                    // columns of the panel still to the right
        if (nin > 0) {
            int rc = gemm_update(c, s, Wb + kk, ld, pr.pd + (size_t)r * pr.ldd + r + kNb, pr.ldd,
                                 tc ? sb.la_hi + (size_t)row_lo * NB + r : nullptr, tc ? sb.la_lo + (size_t)row_lo * NB + r : nullptr,
                                 tc ? pr.ub_hi + (size_t)(kk + kNb) * NB + r : nullptr, tc ? pr.ub_lo + (size_t)(kk + kNb) * NB + r : nullptr,
                                 Wb + kk + kNb, m, nin, kNb);
            if (rc) return rc;
        }
    }
    return GAUSS_CUDA_OK;
}

// P3: A22 -= L21 * U12 for the local rows [row_lo, row_hi) and the columns [col_lo, col_lo + ncols) right of the
// panel (ncols < 0: all of them).  Timed.  max_tiles > 0 bounds how many tiles a CTA of the tile engine takes
// before it exits, so that kernels of the (higher-priority) panel stream get SMs while this launch runs.
static int trailing_update(gauss_cuda_ctx* c, int K, const PanelRef& pr, const StepBufs& sb, int row_lo, int row_hi,
                           int col_lo = -1, int ncols = -1, int max_tiles = 0, int reserve = 0)
{
    cudaStream_t s = sb.s;
    const int ld = c->ld, NB = c->NB, np = c->np;
    const bool tc = c->mode == GAUSS_CUDA_MODE_3XTF32;
    const int k0 = K * NB, kb = blk_rows(c, K);
    if (col_lo < 0) col_lo = k0 + kb;
    if (ncols < 0) ncols = ld - col_lo;
    const int m = row_hi - row_lo;
    if (m <= 0 || ncols <= 0) return GAUSS_CUDA_OK;
    float* Wb = c->W + (size_t)row_lo * ld;
    const int ui = c->upd_count;
    const bool timed = (size_t)(2 * ui + 1) < c->upd_ev.size();
    if (timed) CK(rec_event(c, c->upd_ev[2 * ui], s));
    int rc = gemm_update(c, s, Wb + k0, ld, pr.pu + (col_lo - (k0 + kb)), pr.ldu,
                         tc ? sb.la_hi + (size_t)row_lo * NB : nullptr, tc ? sb.la_lo + (size_t)row_lo * NB : nullptr,
                         tc ? pr.ub_hi + (size_t)col_lo * NB : nullptr, tc ? pr.ub_lo + (size_t)col_lo * NB : nullptr,
                         Wb + col_lo, m, ncols, kb, max_tiles, reserve);
    if (rc) return rc;
    if (timed) {
        CK(rec_event(c, c->upd_ev[2 * ui + 1], s));
        const int real_cols = std::min(col_lo + ncols, np + 1) - col_lo;          // the zero padding is not credited
        c->upd_flops[ui] = 2.0 * (double)m * (double)std::max(real_cols, 0) * (double)kb;
        c->upd_count = ui + 1;
    }
    return GAUSS_CUDA_OK;
}

// Single GPU: P1 -> P2 -> P3 per outer block.  P2 of inner block j (L-col solve of the rows below the panel and
// their in-panel update) only needs the diagonal factor and U-row of block j, so it is forked onto a side stream
// as soon as those are out and runs beside the rest of P1 (whose chain of small kernels -- diagonal factor, panel
// L-col, panel update -- leaves most of the GPU idle); both streams join before P3.
static int outer_step(gauss_cuda_ctx* c, int K, int row_hi, float* ubh, float* ubl)
{
    cudaStream_t s = c->s_main, ss = c->s_side, sw = c->s_wide;
    // plain launches: the panel chain's kernels release their successors early (measured N=2048: 0.99 -> 0.95 ms); under
    // graph capture the replay already has no launch gaps and the attribute bought nothing (0.90 ms either way)
    PdlScope pdl_on(c->pdl && !c->capturing);
    const int ld = c->ld, NB = c->NB, np = c->np;
    const bool tc = c->mode == GAUSS_CUDA_MODE_3XTF32;
    float* lah = c->la_hi[0]; float* lal = c->la_lo[0];
    const StepBufs sb{s, lah, lal};
    const int k0 = K * NB, kb = blk_rows(c, K), ninner = kb / kNb;
    float* P = c->W + (size_t)k0 * ld;
    const int l0 = k0 + kb, m = row_hi - l0;                   // rows below the panel (of those present)
    // (below ~4096 the two-branch CUDA graph costs more launch latency, ~0.2 ms, than the fork saves)
    const bool fork = ss != nullptr && m > 0 && ninner <= 8 && np >= 4096;
    // P1 split at the right edge of the panel: the next diagonal factor only needs the in-panel part of the U-row
    // solve and of the rank-128 update (a few CTAs each), so that chain stays on the main stream; the part right of
    // the panel -- up to two waves of U-row CTAs and a full-width update per inner block at N = 32768 -- follows
    // on its own stream, one inner block behind
    const bool split = sw != nullptr && ninner >= 2 && ninner <= 8 && np >= 8192;
    const int csplit = k0 + kb;                                // first column right of the panel
    float* Wb = c->W + (size_t)l0 * ld;
    int rc;
    bool lcol_done = false;
    for (int j = 0; j < ninner; ++j) {
        const int kk = k0 + j * kNb, r = j * kNb;
        float* D = P + (size_t)r * ld + kk;                    // diagonal block of inner block j
        launch_diag_factor(D, ld, kk, c->status, s);
        const int mrem = kb - (r + kNb);
        // the U-row solve shares its launch with the L-col solve that follows on this stream: the panel rows below
        // (two-level blocking) or, with a single inner block and no fork, all the rows below
        const bool lcol_all = ninner == 1 && !fork && m > 0;
        const int ucols = split ? csplit - (kk + kNb) : np + 1 - (kk + kNb);
        launch_urow_lcol(D, ld,
                         D + kNb, ld, ucols,
                         tc ? ubh + (size_t)(kk + kNb) * NB : nullptr, tc ? ubl + (size_t)(kk + kNb) * NB : nullptr, NB, r,
                         lcol_all ? Wb + kk : P + (size_t)(r + kNb) * ld + kk, ld, lcol_all ? m : mrem,
                         tc ? lah + (size_t)(lcol_all ? l0 : k0 + r + kNb) * NB : nullptr,
                         tc ? lal + (size_t)(lcol_all ? l0 : k0 + r + kNb) * NB : nullptr, NB, r, s);
        lcol_done = lcol_all;
        if (fork || split) CK(cudaEventRecord(c->ev_inner[j], s));
        if (split) {                                           // right of the panel: U-row solve, then the update of the panel rows below
            CK(cudaStreamWaitEvent(sw, c->ev_inner[j], 0));
            launch_urow(D, ld, P + (size_t)r * ld + csplit, ld, np + 1 - csplit,
                        tc ? ubh + (size_t)csplit * NB : nullptr, tc ? ubl + (size_t)csplit * NB : nullptr, NB, r, sw);
            if (mrem > 0) {
                rc = gemm_update(c, sw, P + (size_t)(r + kNb) * ld + kk, ld, P + (size_t)r * ld + csplit, ld,
                                 tc ? lah + (size_t)(k0 + r + kNb) * NB + r : nullptr, tc ? lal + (size_t)(k0 + r + kNb) * NB + r : nullptr,
                                 tc ? ubh + (size_t)csplit * NB + r : nullptr, tc ? ubl + (size_t)csplit * NB + r : nullptr,
                                 P + (size_t)(r + kNb) * ld + csplit, mrem, ld - csplit, kNb);
                if (rc) return rc;
            }
        }
        if (fork) {                                            // P2(j) beside the rest of P1
            CK(cudaStreamWaitEvent(ss, c->ev_inner[j], 0));
            launch_lcol(D, ld, Wb + kk, ld, m,
                        tc ? lah + (size_t)l0 * NB : nullptr, tc ? lal + (size_t)l0 * NB : nullptr, NB, r, ss);
            const int nin = kb - (r + kNb);
            if (nin > 0) {
                rc = gemm_update(c, ss, Wb + kk, ld, D + kNb, ld,
                                 tc ? lah + (size_t)l0 * NB + r : nullptr, tc ? lal + (size_t)l0 * NB + r : nullptr,
                                 tc ? ubh + (size_t)(kk + kNb) * NB + r : nullptr, tc ? ubl + (size_t)(kk + kNb) * NB + r : nullptr,
                                 Wb + kk + kNb, m, nin, kNb);
                if (rc) return rc;
            }
        }
        if (mrem > 0) {                                        // rank-128 update of the panel rows below (in-panel columns only when split)
            const int gcols = split ? csplit - (kk + kNb) : ld - (kk + kNb);
            rc = gemm_update(c, s, P + (size_t)(r + kNb) * ld + kk, ld, D + kNb, ld,
                             tc ? lah + (size_t)(k0 + r + kNb) * NB + r : nullptr, tc ? lal + (size_t)(k0 + r + kNb) * NB + r : nullptr,
                             tc ? ubh + (size_t)(kk + kNb) * NB + r : nullptr, tc ? ubl + (size_t)(kk + kNb) * NB + r : nullptr,
                             P + (size_t)(r + kNb) * ld + kk + kNb, mrem, gcols, kNb);
            if (rc) return rc;
        }
    }
    const float* P0 = P + k0;
    PanelRef pr{P0, ld, P0 + kb, ld, ubh, ubl};
    if (split) {
        CK(cudaEventRecord(c->ev_wide, sw));
        CK(cudaStreamWaitEvent(s, c->ev_wide, 0));
    }
    if (fork) {
        CK(cudaEventRecord(c->ev_side, ss));
        CK(cudaStreamWaitEvent(s, c->ev_side, 0));
    } else if (!lcol_done && (rc = col_panel(c, K, pr, l0, sb, row_hi))) return rc;
    return trailing_update(c, K, pr, sb, l0, row_hi);
}

static int factor_single(gauss_cuda_ctx* c)
{
    for (int K = 0; K < c->nblk; ++K) {
        int rc = outer_step(c, K, c->rows_local, c->ub_hi[0], c->ub_lo[0]);
        if (rc) return rc;
    }
    return GAUSS_CUDA_OK;
}

// Single GPU, host-resident input: elimination overlapped with the upload.  The rows arrive in chunks of whole
// outer blocks on a copy stream.  A chunk is first brought up to date with every panel factored before it arrived
// (L-col solve + trailing update restricted to the chunk's rows: the up-looking half), then its own outer blocks
// are factored right-looking with the rows below limited to the chunk.  Every element receives the same rank-NB
// updates in the same order through the same tiles as in factor_single, so the result is bit-identical; what
// changes is that only the first chunk's transfer is exposed.  The split U operand of every panel stays resident
// (2 x ld x np floats) because later chunks need all of them.
static int factor_streamed(gauss_cuda_ctx* c, const float* A, const float* B)
{
    cudaStream_t s = c->s_main, sc = c->s_copy;
    const int NB = c->NB, ld = c->ld, n = c->n;
    const bool tc = c->mode == GAUSS_CUDA_MODE_3XTF32;
    const size_t ubsz = (size_t)ld * NB;
    // Chunk boundaries in outer blocks.  Early rows carry little work (row r receives r panels' updates) and every
    // chunk pays a latency-bound L-col chain per panel factored before it, so the chunks start small -- little
    // transfer exposed, few earlier panels -- and grow towards the end, where the elimination is far behind the
    // copy stream anyway.  (N = 32768 on B200, measured: 4 equal chunks 160 ms, 8 equal 163 ms, 16 equal 184 ms.)
    std::vector<int> bound;
    if (c->stream_chunks > 0) {
        const int nchunks = std::min(c->stream_chunks, c->nblk), cblk = (c->nblk + nchunks - 1) / nchunks;
        for (int b = cblk; b < c->nblk; b += cblk) bound.push_back(b);
    } else {
        static const int frac32[] = {1, 3, 5, 7, 9, 11, 14, 20};
        for (int f : frac32) {
            const int b = (f * c->nblk + 16) / 32;
            if (b > 0 && b < c->nblk && (bound.empty() || b > bound.back())) bound.push_back(b);
        }
    }
    bound.push_back(c->nblk);
    const int nchunks = (int)bound.size();
    const StepBufs sb{s, c->la_hi[0], c->la_lo[0]};
    CK(cudaMemsetAsync(c->status, 0, sizeof(int), s));
    c->bcast_bytes = 0.0;
    c->upd_count = 0;
    CK(cudaEventRecord(c->ev_gate, s));                        // the copy stream must not overtake earlier work on W
    CK(cudaStreamWaitEvent(sc, c->ev_gate, 0));
    CK(cudaMemcpyAsync(c->hvec, B, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, sc));
    // chunk ci+1 is handed to the copy stream before the kernels of chunk ci are launched: with pinned host memory
    // all of it is asynchronous anyway; with pageable memory the host blocks in the copy while the GPU works
    auto send_chunk = [&](int ci) -> int {
        if (ci >= nchunks) return GAUSS_CUDA_OK;
        const int R0 = ci ? bound[ci - 1] : 0, R1 = bound[ci];
        int rc = upload_blocks(c, c->W, A, R0, R1, sc);
        if (rc) return rc;
        CK(cudaEventRecord(c->ev_chunk[ci], sc));
        if (R1 == c->nblk) CK(cudaEventRecord(c->ev_h2d_done, sc));
        return GAUSS_CUDA_OK;
    };
    { int rc = send_chunk(0); if (rc) return rc; }
    for (int ci = 0; ci < nchunks; ++ci) {
        const int R0 = ci ? bound[ci - 1] : 0, R1 = bound[ci];
        const int r0 = R0 * NB, r1 = std::min(R1 * NB, c->np);
        { int rc = send_chunk(ci + 1); if (rc) return rc; }
        CK(cudaStreamWaitEvent(s, c->ev_chunk[ci], 0));
        int rc;
        for (int K = 0; K < R0; ++K) {                         // panels factored before this chunk arrived
            const int k0 = K * NB, kb = blk_rows(c, K);
            const float* P0 = c->W + (size_t)k0 * ld + k0;
            PanelRef pr{P0, ld, P0 + kb, ld, tc ? c->ub_all_hi + K * ubsz : nullptr, tc ? c->ub_all_lo + K * ubsz : nullptr};
            if ((rc = col_panel(c, K, pr, r0, sb, r1))) return rc;
            if ((rc = trailing_update(c, K, pr, sb, r0, r1))) return rc;
        }
        for (int K = R0; K < R1; ++K)
            if ((rc = outer_step(c, K, r1, tc ? c->ub_all_hi + K * ubsz : nullptr, tc ? c->ub_all_lo + K * ubsz : nullptr))) return rc;
    }
    CK(cudaGetLastError());
    c->factored = true;
    return GAUSS_CUDA_OK;
}

// Multi-GPU with look-ahead.  While the bulk of the trailing update of step K runs on the main stream, the
// whole panel path of step K+1 runs beside it:
//   owner of K+1   row strip of P3(K) (the rows of block K+1) -> P1(K+1) on the panel stream -> packed panel
//                  broadcast on the comm stream (diagonal region first, then U12: the reference's per-step
//                  MPI_Bcast, gauss_mpi :137)
//   every rank     column strip of P3(K) (the columns of block K+1) -> P2(K+1) on the side stream as soon as the
//                  diagonal region has arrived
// so that step K+1 finds its L21 ready and only waits for the U12 region.  Panel buffers and the split L21 / U12
// operands are double buffered by the parity of K; the bulk update's CTAs give up their SM every few tiles so
// the small kernels of the other streams get in.
static int factor_lookahead(gauss_cuda_ctx* c)
{
    cudaStream_t s = c->s_main, sc = c->s_comm;
    cudaStream_t sp = c->s_panel ? c->s_panel : s;
    cudaStream_t ss = c->s_side ? c->s_side : s;
    const int ld = c->ld, NB = c->NB;
    const bool tc = c->mode == GAUSS_CUDA_MODE_3XTF32;
    auto owner_of = [&](int K) { return K % c->world; };
    auto panel_ref = [&](int K) -> PanelRef {
        const int b = K & 1, k0 = K * NB, kb = blk_rows(c, K);
        if (owner_of(K) == c->rank) {                             // the owner reads its panel in place
            const float* P0 = c->W + (size_t)blk_local_row(c, K) * ld + k0;
            return PanelRef{P0, ld, P0 + kb, ld, c->ub_hi[b], c->ub_lo[b]};
        }
        return PanelRef{c->pbuf[b], kb, c->pbuf[b] + (size_t)kb * kb, ld - k0 - kb, c->ub_hi[b], c->ub_lo[b]};
    };
    // enqueue on the comm stream: (owner) pack panel K, then broadcast it into pbuf[K&1] in two pieces --
    // the kb x kb diagonal region first (all that P2 needs), then the U12 region (needed by P3)
    auto enqueue_bcast = [&](int K) -> int {
        const int b = K & 1, k0 = K * NB, kb = blk_rows(c, K), ldu = ld - k0 - kb;
        float* bufd = c->pbuf[b];
        float* bufu = c->pbuf[b] + (size_t)kb * kb;
        CK(cudaStreamWaitEvent(sc, c->ev_free[b], 0));            // last readers of pbuf[b] are done
        const float* P = c->W + (size_t)blk_local_row(c, K) * ld + k0;
        if (owner_of(K) == c->rank) {
            CK(cudaStreamWaitEvent(sc, c->ev_rdiag[b], 0));       // the in-panel chain of P1(K) finished: the diagonal region is final
            CK(cudaMemcpy2DAsync(bufd, (size_t)kb * sizeof(float), P, (size_t)ld * sizeof(float),
                                 (size_t)kb * sizeof(float), kb, cudaMemcpyDeviceToDevice, sc));
        }
        NK(g_nccl.Broadcast(bufd, bufd, (size_t)kb * kb, ncclFloat, owner_of(K), c->comm, sc));
        CK(cudaEventRecord(c->ev_diag[b], sc));
        if (owner_of(K) == c->rank) {
            CK(cudaStreamWaitEvent(sc, c->ev_ready[b], 0));       // all of P1(K) finished
            CK(cudaMemcpy2DAsync(bufu, (size_t)ldu * sizeof(float), P + kb, (size_t)ld * sizeof(float),
                                 (size_t)ldu * sizeof(float), kb, cudaMemcpyDeviceToDevice, sc));
        }
        NK(g_nccl.Broadcast(bufu, bufu, (size_t)kb * ldu, ncclFloat, owner_of(K), c->comm, sc));
        if (owner_of(K) != c->rank) c->bcast_bytes += (double)kb * (kb + ldu) * sizeof(float);
        CK(cudaEventRecord(c->ev_done[b], sc));
        return GAUSS_CUDA_OK;
    };
    // P2(K) on stream `st`: wait for the diagonal region (owner: its own factorisation), then the L-col chain
    auto column_panel_on = [&](int K, cudaStream_t st) -> int {
        const int b = K & 1;
        const PanelRef pr = panel_ref(K);
        if (owner_of(K) == c->rank) {
            CK(cudaStreamWaitEvent(st, c->ev_rdiag[b], 0));       // the owner's P2 needs the diagonal region only
        } else {
            CK(cudaStreamWaitEvent(st, c->ev_diag[b], 0));
            if (tc) rebuild_split_u(c, K, pr, true, st);
        }
        const StepBufs sbk{st, c->la_hi[b], c->la_lo[b]};
        return col_panel(c, K, pr, first_local_block_after(c, K) * NB, sbk);
    };
    int rc;
    CK(cudaEventRecord(c->ev_free[0], s));                        // both panel buffers start out free
    CK(cudaEventRecord(c->ev_free[1], s));
    if (owner_of(0) == c->rank) {
        const StepBufs sb0{s, c->la_hi[0], c->la_lo[0]};
        if ((rc = panel_factor(c, 0, c->ub_hi[0], c->ub_lo[0], sb0, c->s_wide, c->ev_rdiag[0]))) return rc;
        CK(cudaEventRecord(c->ev_ready[0], s));
    }
    if ((rc = enqueue_bcast(0))) return rc;
    if ((rc = column_panel_on(0, s))) return rc;                  // nothing to overlap with yet
    for (int K = 0; K < c->nblk; ++K) {
        const int b = K & 1, k0 = K * NB, kb = blk_rows(c, K);
        const bool mine = owner_of(K) == c->rank;
        const PanelRef pr = panel_ref(K);
        const StepBufs sbm{s, c->la_hi[b], c->la_lo[b]};
        auto mark = [&](int i) { if (c->trace) cudaEventRecord(c->tr_ev[8 * (size_t)K + i], s); };
        mark(0);                                                  // step begins (previous bulk + P2 join done)
        if (!mine) {
            CK(cudaStreamWaitEvent(s, c->ev_done[b], 0));         // the U12 region has arrived
            mark(1);
            if (tc) rebuild_split_u(c, K, pr, false, s);
        } else {
            CK(cudaStreamWaitEvent(s, c->ev_ready[b], 0));        // all of P1(K), the part right of the panel included
            mark(1);
        }
        mark(2);
        const int l0 = first_local_block_after(c, K) * NB;
        if (K + 1 >= c->nblk) {                                   // last block: nothing below it
            CK(cudaEventRecord(c->ev_free[b], s));
            break;
        }
        const int rows_next = blk_rows(c, K + 1), cnext = k0 + kb;
        const bool own_next = owner_of(K + 1) == c->rank;
        int l1 = l0;
        if (own_next) {
            // the rows of block K+1 first, then its factorisation on the panel stream
            if ((rc = trailing_update(c, K, pr, sbm, l0, l0 + rows_next))) return rc;
            if (sp != s) {
                CK(cudaEventRecord(c->ev_fork[b], s));
                CK(cudaStreamWaitEvent(sp, c->ev_fork[b], 0));
            }
            const StepBufs sbp{sp, c->la_hi[b ^ 1], c->la_lo[b ^ 1]};
            if ((rc = panel_factor(c, K + 1, c->ub_hi[b ^ 1], c->ub_lo[b ^ 1], sbp, sp != s ? c->s_wide : nullptr, c->ev_rdiag[b ^ 1]))) return rc;
            CK(cudaEventRecord(c->ev_ready[b ^ 1], sp));
            l1 = l0 + rows_next;
        }
        mark(3);                                                  // row strip issued (owner of K+1 only)
        if ((rc = enqueue_bcast(K + 1))) return rc;
        // the columns of block K+1 for the rows below it, then P2(K+1) on the side stream
        if ((rc = trailing_update(c, K, pr, sbm, l1, c->rows_local, cnext, rows_next))) return rc;
        if (ss != s) {
            CK(cudaEventRecord(c->ev_cols[b], s));
            CK(cudaStreamWaitEvent(ss, c->ev_cols[b], 0));
        }
        mark(4);                                                  // column strip done
        if ((rc = column_panel_on(K + 1, ss))) return rc;
        if (ss != s) CK(cudaEventRecord(c->ev_p2[b ^ 1], ss));
        // the bulk of the update beside all that
        const bool beside = sp != s || ss != s;
        if ((rc = trailing_update(c, K, pr, sbm, l1, c->rows_local, cnext + rows_next, -1, beside ? c->overlap_max_tiles : 0))) return rc;
        CK(cudaEventRecord(c->ev_free[b], s));                    // pbuf[b] / ub[b] / la[b] may be overwritten from here on
        mark(5);                                                  // bulk done
        if (ss != s) CK(cudaStreamWaitEvent(s, c->ev_p2[b ^ 1], 0));
    }
    CK(cudaEventRecord(c->ev_done[0], sc));
    CK(cudaStreamWaitEvent(s, c->ev_done[0], 0));                 // join the comm stream
    return GAUSS_CUDA_OK;
}


// ------------------------------------------------------------------------------------------
// Critical-path schedule.  The chain that bounds a multi-GPU factorisation is
//     diagonal region of block K  ->  what block K+1's owner needs from it  ->  diagonal region of block K+1,
// everything else (the wide U12, the bulk of the trailing update) has a whole step of slack.  This schedule keeps that
// chain as short as the kernels allow:
//   panel(K), owner     E = [[A11, I], [I, 0]] -> in-panel chain on E -> L11|U11, inv(L11), inv(U11)           (lu_cp.cu)
//                       crit(K) = { inv(U11)^T split, U[K, K+1]^T split }  computed first and broadcast first
//                       U12^T = A12^T inv(L11)^T as ONE GEMM into ut[] (broadcast second), U12 itself into the matrix
//   every rank          L21 = A21 inv(U11) as ONE GEMM (+ split), the columns of block K+1 updated at once from crit(K)
//                       (on the owner of K+1 that completes its next diagonal region), then the next A21 split;
//                       ut[] split into the B operand when it arrives; bulk update on the main stream
// The broadcasts are the reference's per-step MPI_Bcast of the pivot row (gauss_mpi/gauss_internal_input.c:137),
// the row ownership its row blocks (:140-155).  world == 1 runs the same code without the collectives.
// ------------------------------------------------------------------------------------------
static int gemm_tc(gauss_cuda_ctx* c, cudaStream_t s, const float* Ahi, const float* Alo, int lda, const float* Bhi, const float* Blo, int ldb,
                   float* C, int ldc, int m, int n, int k, int sign, int beta, int max_tiles = 0)
{
    if (m <= 0 || n <= 0 || k <= 0) return GAUSS_CUDA_OK;
    tc_set_sched_slots(c->sched);
    tc_set_max_tiles(max_tiles);
    tc_set_pair_ok(c->world == 1 || c->dist_pair);
    int rc = tc_gemm_nt(Ahi, Alo, lda, Bhi, Blo, ldb, C, ldc, m, n, k, sign, beta, s);
    tc_set_max_tiles(0);
    tc_set_sched_slots(nullptr);
    if (rc) return cuda_fail((cudaError_t)rc, "tc_gemm_nt", __LINE__);
    return GAUSS_CUDA_OK;
}

// In-panel chain on the extended scratch: eliminates the first kb pivots of E (2kb x 2kb, leading dimension lde).
static int ext_chain(gauss_cuda_ctx* c, int k0, int kb, cudaStream_t s)
{
    const int NB = c->NB, lde = 2 * NB, n2 = 2 * kb, ninner = kb / kNb;
    float* E = c->ext;
    for (int j = 0; j < ninner; ++j) {
        const int r = j * kNb, rest = n2 - (r + kNb);
        float* D = E + (size_t)r * lde + r;
        launch_diag_factor(D, lde, k0 + r, c->status, s);
        launch_urow_lcol(D, lde,
                         D + kNb, lde, rest, c->ext_ub[0] + (size_t)(r + kNb) * NB, c->ext_ub[1] + (size_t)(r + kNb) * NB, NB, r,
                         E + (size_t)(r + kNb) * lde + r, lde, rest, c->ext_la[0] + (size_t)(r + kNb) * NB, c->ext_la[1] + (size_t)(r + kNb) * NB, NB, r, s);
        // rank-128 update of what is left of E; the last inner block only has the two identity strips left, whose
        // product lands in the unused bottom-right block, so it is skipped
        if (j + 1 < ninner) {
            int rc = gemm_tc(c, s, c->ext_la[0] + (size_t)(r + kNb) * NB + r, c->ext_la[1] + (size_t)(r + kNb) * NB + r, NB,
                             c->ext_ub[0] + (size_t)(r + kNb) * NB + r, c->ext_ub[1] + (size_t)(r + kNb) * NB + r, NB,
                             E + (size_t)(r + kNb) * lde + r + kNb, lde, rest, rest, kNb, -1, 1);
            if (rc) return rc;
        }
    }
    return GAUSS_CUDA_OK;
}

static int factor_cp(gauss_cuda_ctx* c)
{
    cudaStream_t s = c->s_main, sc = c->s_comm, sp = c->s_panel, ss = c->s_side, sw = c->s_wide;
    if (!sp || !ss || !sw || !sc) { snprintf(g_err, sizeof g_err, "critical-path schedule needs its side streams (GAUSS_CUDA_DIST_FORK/WIDE_SPLIT must stay on)"); return GAUSS_CUDA_BAD_ARG; }
    const int ld = c->ld, NB = c->NB, nblk = c->nblk, lde = 2 * NB;
    const size_t nb2 = (size_t)NB * NB;
    auto owner_of = [&](int K) { return K % c->world; };
    auto iut_hi = [&](int b) { return c->crit[b]; };
    auto iut_lo = [&](int b) { return c->crit[b] + nb2; };
    auto uc_hi = [&](int b) { return c->crit[b] + 2 * nb2; };
    auto uc_lo = [&](int b) { return c->crit[b] + 3 * nb2; };
    // columns right of block K: [crit block K+1 | near: the next cp_near blocks | far: the rest]; near_end = first far column
    auto near_end = [&](int K) { return std::min(ld, (K + 2 + c->cp_near) * NB); };
    int rc;
    // peer push: the owner of K stores crit(K) into the box of the owner of K+1 (always the next rank); the receiver
    // double-buffers by the parity of its own use count ((K+1) / world) and waits for flag words 0 (inv(U11)^T) and 32 (U^T)
    c->solve_seq += 1;
    const bool p2p = c->p2p && c->world > 1;
    auto pq = [&](int K) { return ((K + 1) / c->world) & 1; };
    auto seq_of = [&](int K) { return (unsigned)(c->solve_seq * (unsigned)nblk + (unsigned)K + 1u); };
    auto box_crit = [&](float* box, int q) { return box + 256 + (size_t)q * 4 * nb2; };
    auto wait_flag = [&](cudaStream_t st, int word, unsigned value) -> int {
        const unsigned* f = reinterpret_cast<const unsigned*>(c->pbox) + word;
        if (c->memops) {
            if (g_wait_value32(st, (unsigned long long)(uintptr_t)f, value, 0x0 /* CU_STREAM_WAIT_VALUE_GEQ */) != 0) {
                snprintf(g_err, sizeof g_err, "cuStreamWaitValue32 failed"); return GAUSS_CUDA_CUDA_ERR;
            }
        } else launch_wait_flag(f, value, st);
        return GAUSS_CUDA_OK;
    };

    // panel(K) on its owner, first half: diagonal region -> L11|U11, inv(L11) (il), inv(U11)^T (crit[b].iut)
    auto panel_diag = [&](int K) -> int {
        const int b = K & 1, k0 = K * NB, kb = blk_rows(c, K);
        float* A11 = c->W + (size_t)blk_local_row(c, K) * ld + k0;
        PdlScope pdl_on(c->pdl);                                    // the chain's kernels release each other early
        CK(cudaStreamWaitEvent(sp, c->ev_critfree[b], 0));        // crit[b] (iut) of step K-2 has been consumed here
        launch_ext_prepare(A11, ld, c->ext, lde, kb, sp);
        if ((rc = ext_chain(c, k0, kb, sp))) return rc;
        CK(cudaStreamWaitEvent(sp, c->ev_wjoin, 0));              // il is single-buffered: the previous row panel's GEMMs (one rank: ours) have read it
        const bool push = p2p && K + 1 < nblk;
        float* pc = push ? box_crit(c->peer_pbox, pq(K)) : nullptr;
        launch_ext_finish(c->ext, lde, kb, A11, ld, c->il[0], c->il[1], iut_hi(b), iut_lo(b), NB,
                          pc, pc ? pc + nb2 : nullptr, c->push_counters, push ? reinterpret_cast<unsigned*>(c->peer_pbox) : nullptr, seq_of(K),
                          c->iu_all ? c->iu_all + (size_t)(K / c->world) * nb2 : nullptr, sp);
        CK(cudaEventRecord(c->ev_inv[b], sp));
        return GAUSS_CUDA_OK;
    };
    // panel(K) on its owner, second half (the row panel right of the diagonal region must be final: ev_rowrest):
    //   crit columns (those of block K+1) first on the panel stream, the rest on the wide stream
    auto panel_row = [&](int K, bool have_rowrest) -> int {
        const int b = K & 1, k0 = K * NB, kb = blk_rows(c, K);
        const int c1 = k0 + kb;                                     // first column right of the diagonal region
        const int ncols = ld - c1;
        float* P = c->W + (size_t)blk_local_row(c, K) * ld;       // the row panel (kb rows)
        const int ccrit = (K + 1 < nblk) ? blk_rows(c, K + 1) : 0;  // columns of block K+1
        CK(cudaStreamWaitEvent(sp, c->ev_wjoin, 0));              // a12t is single-buffered too (previous row panel on the wide stream)
        if (have_rowrest) {                                        // the crit columns only need the near part of the previous step
            CK(cudaStreamWaitEvent(sp, c->ev_rownear[b ^ 1], 0));
            CK(cudaStreamWaitEvent(sw, c->ev_rownear[b ^ 1], 0));
            CK(cudaStreamWaitEvent(sw, c->ev_rowrest[b ^ 1], 0));
        }
        if (ccrit > 0) {
            PdlScope pdl_on(c->pdl);
            CK(cudaStreamWaitEvent(sp, c->ev_utfree[b], 0));        // ut[b] of step K-2 has been split everywhere it is read here
            launch_split_transpose(P + c1, ld, c->a12t[0] + (size_t)c1 * NB, c->a12t[1] + (size_t)c1 * NB, NB, kb, ccrit, kb, sp);
            CK(cudaEventRecord(c->ev_p4a, sp));
            // U[K, K+1]^T = A12^T inv(L11)^T  (kb_next x kb), raw into ut[b], then split into the crit package
            if ((rc = gemm_tc(c, sp, c->a12t[0] + (size_t)c1 * NB, c->a12t[1] + (size_t)c1 * NB, NB, c->il[0], c->il[1], NB,
                              c->ut[b] + (size_t)c1 * NB, NB, ccrit, kb, kb, +1, 0))) return rc;
            float* pc = p2p ? box_crit(c->peer_pbox, pq(K)) : nullptr;
            launch_crit_split(c->ut[b] + (size_t)c1 * NB, NB, uc_hi(b), uc_lo(b), NB, ccrit, kb,
                              pc ? pc + 2 * nb2 : nullptr, pc ? pc + 3 * nb2 : nullptr, c->push_counters + 8,
                              p2p ? reinterpret_cast<unsigned*>(c->peer_pbox) + 32 : nullptr, seq_of(K), sp);
            CK(cudaEventRecord(c->ev_critlocal[b], sp));
            if (c->trace && K > 0) cudaEventRecord(c->tr_ev[8 * (size_t)(K - 1) + 4], sp);
        } else {
            CK(cudaEventRecord(c->ev_p4a, sp));
        }
        // the rest of the row panel on the wide stream
        CK(cudaStreamWaitEvent(sw, c->ev_inv[b], 0));
        CK(cudaStreamWaitEvent(sw, c->ev_utfree[b], 0));
        const int crest = c1 + ccrit, cn = std::max(crest, near_end(K));
        for (int part = 0; part < 2; ++part) {                     // near columns, then far columns
            const int ca = part ? cn : crest, cb = part ? ld : cn;
            if (cb > ca) {
                launch_split_transpose(P + ca, ld, c->a12t[0] + (size_t)ca * NB, c->a12t[1] + (size_t)ca * NB, NB, kb, cb - ca, kb, sw);
                if (K + 1 < nblk) {
                    if ((rc = gemm_tc(c, sw, c->a12t[0] + (size_t)ca * NB, c->a12t[1] + (size_t)ca * NB, NB, c->il[0], c->il[1], NB,
                                      c->ut[b] + (size_t)ca * NB, NB, cb - ca, kb, kb, +1, 0))) return rc;
                }
            }
            if (part == 0) {
                CK(cudaStreamWaitEvent(sw, c->ev_critlocal[b], 0)); // (the crit rows of ut[b] travel with the near payload)
                CK(cudaEventRecord(c->ev_nearlocal[b], sw));
            }
        }
        CK(cudaEventRecord(c->ev_utlocal[b], sw));
        // U12 itself, in place, for back substitution (solveGauss reads the owner's rows): inv(L11) A12
        CK(cudaStreamWaitEvent(sw, c->ev_p4a, 0));
        if ((rc = gemm_tc(c, sw, c->il[0], c->il[1], NB, c->a12t[0] + (size_t)c1 * NB, c->a12t[1] + (size_t)c1 * NB, NB,
                          P + c1, ld, kb, ncols, kb, +1, 0))) return rc;
        CK(cudaEventRecord(c->ev_wjoin, sw));
        return GAUSS_CUDA_OK;
    };
    // comm(K): crit package, then the bulk payload, from the owner to everyone (no-ops with one rank)
    auto comm = [&](int K) -> int {
        if (K + 1 >= nblk) return GAUSS_CUDA_OK;                   // the last block has nothing below or right of it
        const int b = K & 1, k0 = K * NB, kb = blk_rows(c, K), c1 = k0 + kb;
        const bool mine = owner_of(K) == c->rank;
        if (mine) CK(cudaStreamWaitEvent(sc, c->ev_critlocal[b], 0));
        else      CK(cudaStreamWaitEvent(sc, c->ev_critfree[b], 0));
        if (c->world > 1) {
            NK(g_nccl.Broadcast(c->crit[b], c->crit[b], 4 * nb2, ncclFloat, owner_of(K), c->comm, sc));
            if (!mine) c->bcast_bytes += 4.0 * nb2 * sizeof(float);
        }
        CK(cudaEventRecord(c->ev_crit[b], sc));
        // U12^T: the near columns (crit block included) first, the far columns second
        const int cn = std::max(c1, near_end(K));
        if (mine) CK(cudaStreamWaitEvent(sc, c->ev_nearlocal[b], 0));
        else      CK(cudaStreamWaitEvent(sc, c->ev_utfree[b], 0));
        if (c->world > 1) {
            const size_t cnt = (size_t)(cn - c1) * NB;
            NK(g_nccl.Broadcast(c->ut[b] + (size_t)c1 * NB, c->ut[b] + (size_t)c1 * NB, cnt, ncclFloat, owner_of(K), c->comm, sc));
            if (!mine) c->bcast_bytes += (double)cnt * sizeof(float);
        }
        CK(cudaEventRecord(c->ev_near[b], sc));
        if (mine) CK(cudaStreamWaitEvent(sc, c->ev_utlocal[b], 0));
        if (c->world > 1 && ld > cn) {
            const size_t cnt = (size_t)(ld - cn) * NB;
            NK(g_nccl.Broadcast(c->ut[b] + (size_t)cn * NB, c->ut[b] + (size_t)cn * NB, cnt, ncclFloat, owner_of(K), c->comm, sc));
            if (!mine) c->bcast_bytes += (double)cnt * sizeof(float);
        }
        CK(cudaEventRecord(c->ev_ut[b], sc));
        return GAUSS_CUDA_OK;
    };

    // ---- prologue
    for (int i = 0; i < 2; ++i) {                                  // every double-buffered resource starts out free
        CK(cudaEventRecord(c->ev_critfree[i], s));
        CK(cudaEventRecord(c->ev_utfree[i], s));
        CK(cudaEventRecord(c->ev_free[i], s));
    }
    CK(cudaEventRecord(c->ev_fork[0], s));                          // the side streams start after whatever precedes on the main stream
    for (cudaStream_t st : {sp, ss, sw, sc}) CK(cudaStreamWaitEvent(st, c->ev_fork[0], 0));
    {
        const int l0 = first_local_block_after(c, 0) * NB, m = c->rows_local - l0;
        if (m > 0 && nblk > 1)
            launch_split_rowmajor(c->W + (size_t)l0 * ld, ld, c->as_[0] + (size_t)l0 * NB, c->as_[1] + (size_t)l0 * NB, NB, m, blk_rows(c, 0), ss);
    }
    if (owner_of(0) == c->rank) {
        if ((rc = panel_diag(0))) return rc;
        if ((rc = panel_row(0, false))) return rc;
    }
    if ((rc = comm(0))) return rc;

    // ---- steps
    for (int K = 0; K + 1 < nblk; ++K) {
        const int b = K & 1, k0 = K * NB, kb = blk_rows(c, K);
        const int rows_next = blk_rows(c, K + 1), cnext = k0 + kb;
        const bool own_next = owner_of(K + 1) == c->rank;
        const int l0 = first_local_block_after(c, K) * NB;          // first local row below block K
        const int l1 = own_next ? l0 + rows_next : l0;              // first local row below block K+1

        auto mark = [&](int i, cudaStream_t st) { if (c->trace) cudaEventRecord(c->tr_ev[8 * (size_t)K + i], st); };
        // 1. side stream: L21 = A21 inv(U11) (+ split), columns of block K+1, next A21 split
        // the next owner reads crit(K) from its peer box (pushed by the owner of K), everyone else from the broadcast
        const bool from_box = p2p && own_next;
        const float* c_iut_hi = from_box ? box_crit(c->pbox, pq(K)) : iut_hi(b);
        const float* c_iut_lo = c_iut_hi + nb2;
        const float* c_uc_hi = c_iut_hi + 2 * nb2;
        const float* c_uc_lo = c_iut_hi + 3 * nb2;
        if (from_box) { if ((rc = wait_flag(ss, 0, seq_of(K)))) return rc; }
        else CK(cudaStreamWaitEvent(ss, c->ev_crit[b], 0));
        CK(cudaStreamWaitEvent(ss, c->ev_free[b], 0));              // la[b] of step K-2 no longer read by its bulk update
        mark(0, ss);
        // L21 of rows [ra, rb) (local), then their columns of block K+1
        auto l21_rows = [&](int ra, int rb) -> int {
            if (rb <= ra) return GAUSS_CUDA_OK;
            float* Wr = c->W + (size_t)ra * ld;
            int r2 = gemm_tc(c, ss, c->as_[0] + (size_t)ra * NB, c->as_[1] + (size_t)ra * NB, NB, c_iut_hi, c_iut_lo, NB,
                             Wr + k0, ld, rb - ra, kb, kb, +1, 0);
            if (r2) return r2;
            launch_split_rowmajor(Wr + k0, ld, c->la_hi[b] + (size_t)ra * NB, c->la_lo[b] + (size_t)ra * NB, NB, rb - ra, kb, ss);
            return GAUSS_CUDA_OK;
        };
        auto colstrip_rows = [&](int ra, int rb) -> int {
            if (rb <= ra) return GAUSS_CUDA_OK;
            return gemm_tc(c, ss, c->la_hi[b] + (size_t)ra * NB, c->la_lo[b] + (size_t)ra * NB, NB, c_uc_hi, c_uc_lo, NB,
                           c->W + (size_t)ra * ld + cnext, ld, rb - ra, rows_next, kb, -1, 1);
        };
        // Column bookkeeping.  Panel j reaches block column J through: the column strip (J = j+1, this stream), the NEAR
        // update (j+2 <= J <= j+1+near, this stream, below) or the FAR update (J beyond, main stream).  Block column K+1 is
        // read here (next A21 split, next diagonal region) and block column K+2 by the next row panel's crit columns: with
        // near >= 2 the far updates they depend on are those of steps <= K-2, which ev_free[b] above already covers; all
        // near updates and column strips precede this point in stream order.
        if (own_next) {                                            // the rows of block K+1 first: they complete the next diagonal region
            PdlScope pdl_on(c->pdl);
            if ((rc = l21_rows(l0, l1))) return rc;
            if (from_box && (rc = wait_flag(ss, 32, seq_of(K)))) return rc;
            if ((rc = colstrip_rows(l0, l1))) return rc;
            CK(cudaEventRecord(c->ev_colstrip[b], ss));
            mark(1, ss);
            if ((rc = l21_rows(l1, c->rows_local))) return rc;
            CK(cudaEventRecord(c->ev_la[b], ss));
            if ((rc = colstrip_rows(l1, c->rows_local))) return rc;
        } else {
            if ((rc = l21_rows(l0, c->rows_local))) return rc;
            CK(cudaEventRecord(c->ev_la[b], ss));
            mark(1, ss);
            if ((rc = colstrip_rows(l0, c->rows_local))) return rc;
            CK(cudaEventRecord(c->ev_colstrip[b], ss));
        }
        CK(cudaEventRecord(c->ev_critfree[b], ss));
        mark(2, ss);
        if (c->rows_local - l1 > 0 && K + 2 < nblk)
            launch_split_rowmajor(c->W + (size_t)l1 * ld + cnext, ld, c->as_[0] + (size_t)l1 * NB, c->as_[1] + (size_t)l1 * NB, NB,
                                  c->rows_local - l1, rows_next, ss);
        // near columns of U12^T(K): B operand, then the near update of every local row below block K (the rows of block K+1
        // first on its owner: its row panel's crit columns wait for them)
        const int cn = std::max(cnext, near_end(K)), cfirst = cnext + rows_next;    // first far column, first column right of block K+1
        const PanelRef pr{nullptr, 0, nullptr, 0, c->ub_hi[b], c->ub_lo[b]};
        {
            const StepBufs sbs{ss, c->la_hi[b], c->la_lo[b]};
            CK(cudaStreamWaitEvent(ss, c->ev_near[b], 0));
            launch_split_rowmajor(c->ut[b] + (size_t)cnext * NB, NB, c->ub_hi[b] + (size_t)cnext * NB, c->ub_lo[b] + (size_t)cnext * NB, NB,
                                  cn - cnext, kb, ss);
            CK(cudaEventRecord(c->ev_nearsplit[b], ss));
            if (own_next) {
                if (cn > cfirst && (rc = trailing_update(c, K, pr, sbs, l0, l1, cfirst, cn - cfirst))) return rc;
                CK(cudaEventRecord(c->ev_rownear[b], ss));
            }
            if (cn > cfirst && !c->cp_nobulk && (rc = trailing_update(c, K, pr, sbs, l1, c->rows_local, cfirst, cn - cfirst, 0, c->cp_reserve))) return rc;
        }

        // 2. panel stream (owner of K+1): its diagonal region is final after the column strip
        if (own_next) {
            CK(cudaStreamWaitEvent(sp, c->ev_colstrip[b], 0));
            if ((rc = panel_diag(K + 1))) return rc;
            mark(3, sp);
        }

        // 3. main stream: the far columns of U12^T(K) arrive -> B operand; the far part of block K+1's rows
        const StepBufs sbm{s, c->la_hi[b], c->la_lo[b]};
        CK(cudaStreamWaitEvent(s, c->ev_ut[b], 0));
        mark(6, s);
        if (ld > cn)
            launch_split_rowmajor(c->ut[b] + (size_t)cn * NB, NB, c->ub_hi[b] + (size_t)cn * NB, c->ub_lo[b] + (size_t)cn * NB, NB, ld - cn, kb, s);
        CK(cudaStreamWaitEvent(s, c->ev_nearsplit[b], 0));          // ut[b] is free once both parts have been split
        CK(cudaEventRecord(c->ev_utfree[b], s));
        CK(cudaStreamWaitEvent(s, c->ev_la[b], 0));
        if (own_next) {
            if (ld > std::max(cn, cfirst) && (rc = trailing_update(c, K, pr, sbm, l0, l1, std::max(cn, cfirst), -1))) return rc;
            CK(cudaEventRecord(c->ev_rowrest[b], s));
            // 4. the row panel of block K+1: crit package first, wide part beside it
            if ((rc = panel_row(K + 1, true))) return rc;
        }
        mark(7, s);
        // 5. communication of step K+1
        if ((rc = comm(K + 1))) return rc;
        // 6. bulk of the trailing update (CTAs give up their SM every few tiles: the other streams have priority)
        if (!c->cp_nobulk && ld > std::max(cn, cfirst) &&
            (rc = trailing_update(c, K, pr, sbm, l1, c->rows_local, std::max(cn, cfirst), -1,
                                  c->cp_reserve > 0 ? 0 : c->overlap_max_tiles, c->cp_reserve))) return rc;
        CK(cudaEventRecord(c->ev_free[b], s));
        mark(5, s);
    }
    // ---- join everything back into the main stream
    CK(cudaEventRecord(c->ev_join[0], sp)); CK(cudaStreamWaitEvent(s, c->ev_join[0], 0));
    CK(cudaEventRecord(c->ev_join[1], ss)); CK(cudaStreamWaitEvent(s, c->ev_join[1], 0));
    CK(cudaEventRecord(c->ev_wjoin, sw));   CK(cudaStreamWaitEvent(s, c->ev_wjoin, 0));
    CK(cudaEventRecord(c->ev_done[0], sc)); CK(cudaStreamWaitEvent(s, c->ev_done[0], 0));
    return GAUSS_CUDA_OK;
}

// Partial pivoting (gauss_external_input.c:123-182): single-level blocking, one 128-column panel at a
// time -- pivoted panel factorisation over all rows below, interchanges applied to the right-hand part,
// U-row solve, rank-128 trailing update (the no-pivot kernels from there on).
static int factor_pivot(gauss_cuda_ctx* c)
{
    cudaStream_t s = c->s_main;
    const int ld = c->ld, NB = c->NB, np = c->np;
    const bool tc = c->mode == GAUSS_CUDA_MODE_3XTF32;
    for (int k0 = 0; k0 < np; k0 += kNb) {
        const int below = k0 + kNb;
        launch_panel_pivot(c->W, ld, np, k0, c->pt, np, c->ipiv, c->status,
                           tc ? c->la_hi[0] + (size_t)below * NB : nullptr, tc ? c->la_lo[0] + (size_t)below * NB : nullptr, NB, s);
        float* P = c->W + (size_t)k0 * ld;
        launch_urow(P + k0, ld, P + below, ld, np + 1 - below,
                    tc ? c->ub_hi[0] + (size_t)below * NB : nullptr, tc ? c->ub_lo[0] + (size_t)below * NB : nullptr, NB, 0, s);
        const int m = np - below;
        if (m > 0) {
            float* Wb = c->W + (size_t)below * ld;
            int rc = gemm_update(c, s, Wb + k0, ld, P + below, ld,
                                 tc ? c->la_hi[0] + (size_t)below * NB : nullptr, tc ? c->la_lo[0] + (size_t)below * NB : nullptr,
                                 tc ? c->ub_hi[0] + (size_t)below * NB : nullptr, tc ? c->ub_lo[0] + (size_t)below * NB : nullptr,
                                 Wb + below, m, ld - below, kNb);
            if (rc) return rc;
        }
    }
    return GAUSS_CUDA_OK;
}

static int factor(gauss_cuda_ctx* c)
{
    CK(cudaMemsetAsync(c->status, 0, sizeof(int), c->s_main));
    c->bcast_bytes = 0.0;
    c->upd_count = 0;
    const auto host_t0 = std::chrono::steady_clock::now();
    int rc = c->pivot ? factor_pivot(c)
           : c->cp ? factor_cp(c)
           : c->world > 1 ? factor_lookahead(c)
           : factor_single(c);
    c->host_issue_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - host_t0).count();
    if (rc) return rc;
    CK(cudaGetLastError());
    c->factored = true;
    return GAUSS_CUDA_OK;
}

// Back substitution, solveGauss (gauss_internal_input.c:144-159): outer blocks from the last up;
// the owner solves its diagonal blocks, x_K travels to everyone, every rank folds it into its y.
static int backsub(gauss_cuda_ctx* c)
{
    cudaStream_t s = c->s_main;
    const int ld = c->ld, NB = c->NB, np = c->np;
    if (c->world == 1) {                   // one launch; U is streamed once, W stays untouched
        launch_trsv_fused(c->W, ld, np, np, c->xvec, c->trsv_ctrl, s);
        CK(cudaGetLastError());
        return GAUSS_CUDA_OK;
    }
    if (c->cp && c->p2p && c->iu_all) {
        // peer-memory form: the owner multiplies by the stored inv(U11) and writes x_K into every rank's receive vector
        // itself (one kernel, fused with its broadcast); the others wait on a flag word, then fold x_K into their rows
        const size_t nb2 = (size_t)NB * NB;
        float* xr = c->pbox + 256 + 8 * nb2;                         // this rank's receive vector (np floats)
        c->bs_count += 1;
        for (int K = c->nblk - 1; K >= 0; --K) {
            const int k0 = K * NB, kb = blk_rows(c, K), owner = K % c->world;
            const unsigned seq = c->bs_count * (unsigned)(c->nblk + 1) + (unsigned)(c->nblk - K);
            if (owner == c->rank) {
                BsPeers peers{};
                for (int r = 0; r < c->world; ++r) {
                    if (r == c->rank) continue;
                    peers.x[peers.n] = c->peer_box[r] + 256 + 8 * nb2 + k0;
                    peers.flag[peers.n] = reinterpret_cast<unsigned*>(c->peer_box[r]) + 64;
                    ++peers.n;
                }
                const float* P = c->W + (size_t)blk_local_row(c, K) * ld;
                launch_bs_solve_push(c->iu_all + (size_t)(K / c->world) * nb2, NB, P + np, ld, kb, xr + k0, peers, c->push_counters + 16, seq, s);
            } else {
                const unsigned* f = reinterpret_cast<const unsigned*>(c->pbox) + 64;
                if (c->memops) {
                    if (g_wait_value32(s, (unsigned long long)(uintptr_t)f, seq, 0x0) != 0) { snprintf(g_err, sizeof g_err, "cuStreamWaitValue32 failed"); return GAUSS_CUDA_CUDA_ERR; }
                } else launch_wait_flag(f, seq, s);
                c->bcast_bytes += (double)kb * sizeof(float);
            }
            const int mabove = local_blocks_before(c, K) * NB;
            launch_trsv_gemv_wide(c->W + k0, ld, np - k0, xr + k0, mabove, kb, s);
        }
        CK(cudaMemcpyAsync(c->xvec, xr, (size_t)np * sizeof(float), cudaMemcpyDeviceToDevice, s));
        CK(cudaGetLastError());
        return GAUSS_CUDA_OK;
    }
    for (int K = c->nblk - 1; K >= 0; --K) {
        const int k0 = K * NB, kb = blk_rows(c, K), owner = K % c->world;
        if (owner == c->rank) {                    // the owner's kb x kb diagonal system, one fused launch
            const float* P = c->W + (size_t)blk_local_row(c, K) * ld + k0;
            launch_trsv_fused(P, ld, kb, np - k0, c->xvec + k0, c->trsv_ctrl, s);
        }
        NK(g_nccl.Broadcast(c->xvec + k0, c->xvec + k0, (size_t)kb, ncclFloat, owner, c->comm, s));
        if (owner != c->rank) c->bcast_bytes += (double)kb * sizeof(float);
        const int mabove = local_blocks_before(c, K) * NB;
        launch_trsv_gemv_wide(c->W + k0, ld, np - k0, c->xvec + k0, mabove, kb, s);
    }
    CK(cudaGetLastError());
    return GAUSS_CUDA_OK;
}

static int finish_stats(gauss_cuda_ctx* c, gauss_cuda_stats* st, bool h2d, bool fac, bool bs, bool d2h)
{
    CK(cudaStreamSynchronize(c->s_main));
    int flag = 0;
    CK(cudaMemcpy(&flag, c->status, sizeof(int), cudaMemcpyDeviceToHost));
    gauss_cuda_stats& o = c->last;
    auto el = [&](int a, int b) { float ms = 0.f; cudaEventElapsedTime(&ms, c->ev[a], c->ev[b]); return (double)ms * 1e-3; };
    o.h2d_s = h2d ? el(0, 1) : 0.0;
    if (h2d && c->last_streamed) {             // overlapped schedule: the upload runs beside the elimination
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c->ev[0], c->ev_h2d_done) == cudaSuccess) o.h2d_s = (double)ms * 1e-3;
    }
    o.overlapped_upload = c->last_streamed ? 1 : 0;
    o.factor_s = fac ? el(1, 2) : 0.0;
    o.backsub_s = bs ? el(2, 3) : 0.0;
    o.d2h_s = d2h ? el(3, 4) : 0.0;
    o.bcast_bytes = c->bcast_bytes;
    o.zero_pivot_row = flag ? flag - 1 : -1;
    o.n = c->n; o.n_padded = c->np; o.nb_outer = c->NB; o.nb_inner = kNb; o.mode = c->mode; o.rank = c->rank; o.world = c->world;
    o.kernel_launches = g_launch_count;
    o.graph_replayed = c->last_graph ? 1 : 0;
    c->last_graph = false;
    o.update_s = 0.0; o.update_flops = 0.0; o.update_launches = 0;
    if (fac) {
        for (int i = 0; i < c->upd_count; ++i) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, c->upd_ev[2 * i], c->upd_ev[2 * i + 1]) == cudaSuccess) {
                o.update_s += (double)ms * 1e-3;
                o.update_flops += c->upd_flops[i];
                o.update_launches += 1;
            }
        }
    }
    if (c->trace && fac && c->cp) {
        // critical-path schedule, per rank: side stream [0 crit arrived -> 1 L21 GEMM + split -> 2 column strip], panel stream
        // (owner of K+1) [2 -> 3 diagonal region chain -> 4 crit columns of U12], main stream [6 U12^T arrived -> 7 split +
        // row strip -> 5 bulk done]; with one rank [4 of step K -> 0 of step K+1] is the hand-over
        double a_p2 = 0, a_cs = 0, a_chain = 0, a_cu = 0, a_hand = 0, a_utwait = 0, a_split = 0, a_bulk = 0, a_crit2crit = 0;
        int n_chain = 0;
        auto el2 = [&](int Ka, int ia, int Kb, int ib, double& acc) -> bool {
            if (Ka < 0 || Kb < 0 || Ka >= c->nblk || Kb >= c->nblk) return false;
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, c->tr_ev[8 * (size_t)Ka + ia], c->tr_ev[8 * (size_t)Kb + ib]) != cudaSuccess) { cudaGetLastError(); return false; }
            acc += ms; return true;
        };
        for (int K = 0; K + 1 < c->nblk; ++K) {
            el2(K, 0, K, 1, a_p2); el2(K, 1, K, 2, a_cs);
            if (K % c->world == (c->rank + c->world - 1) % c->world && el2(K, 1, K, 3, a_chain)) ++n_chain;
            if (K % c->world == (c->rank + c->world - 1) % c->world) el2(K, 3, K, 4, a_cu);
            if (c->world == 1) el2(K, 4, K + 1, 0, a_hand);
            el2(K, 0, K + 1, 0, a_crit2crit);
            el2(K - 1, 5, K, 6, a_utwait); el2(K, 6, K, 7, a_split); el2(K, 7, K, 5, a_bulk);
        }
        cudaGetLastError();
        fprintf(stderr, "[gauss_cuda trace cp] rank %d/%d nb %d: side L21 %.2f ms, col strip %.2f | panel chain %.2f (%d panels), crit U %.2f, hand-over %.2f | "
                        "crit-to-crit %.2f | main: wait U12 %.2f, split+row strip %.2f, bulk %.2f (factor %.2f ms, host issue %.2f ms)\n",
                c->rank, c->world, c->NB, a_p2, a_cs, a_chain, n_chain, a_cu, a_hand, a_crit2crit, a_utwait, a_split, a_bulk, o.factor_s * 1e3, c->host_issue_ms);
    } else if (c->trace && fac) {
        // where the main stream's time went: [0->1] waiting for the U12 broadcast, [1->2] rebuilding the split,
        // [2->3] row strip, [3->4] column strip, [4->5] bulk update, [5->next 0] waiting for P2 of the next panel
        double acc[6] = {0, 0, 0, 0, 0, 0};
        for (int K = 0; K + 1 < c->nblk; ++K) {
            for (int i = 0; i < 6; ++i) {
                cudaEvent_t a = c->tr_ev[8 * (size_t)K + i];
                cudaEvent_t b2 = i < 5 ? c->tr_ev[8 * (size_t)K + i + 1] : c->tr_ev[8 * (size_t)(K + 1)];
                float ms = 0.f;
                if (cudaEventElapsedTime(&ms, a, b2) == cudaSuccess) acc[i] += ms;
            }
        }
        cudaGetLastError();
        fprintf(stderr, "[gauss_cuda trace] rank %d/%d: wait_U12 %.2f ms, rebuild %.2f, row_strip %.2f, col_strip %.2f, bulk %.2f, wait_P2 %.2f (factor %.2f ms)\n",
                c->rank, c->world, acc[0], acc[1], acc[2], acc[3], acc[4], acc[5], o.factor_s * 1e3);
    }
    if (st) *st = o;
    return flag ? GAUSS_CUDA_ZERO_PIVOT : GAUSS_CUDA_OK;
}

// X is complete on every rank after back substitution; B after elimination is spread over the ranks.
static int download_results(gauss_cuda_ctx* c, float* X, float* B_out)
{
    cudaStream_t s = c->s_main;
    if (X) CK(cudaMemcpyAsync(X, c->xvec, (size_t)c->n * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (B_out) {
        // hvec holds this rank's rows of y = L^-1 b (snapshot_rhs); the other ranks' rows arrive here,
        // one small broadcast per outer block (collective: every rank must ask for B_out or none).
        if (c->world > 1)
            for (int R = 0; R < c->nblk; ++R)
                NK(g_nccl.Broadcast(c->hvec + R * c->NB, c->hvec + R * c->NB, (size_t)blk_rows(c, R), ncclFloat, R % c->world, c->comm, s));
        CK(cudaMemcpyAsync(B_out, c->hvec, (size_t)c->n * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    return GAUSS_CUDA_OK;
}

// After elimination, snapshot this rank's rows of the RHS column (y = L^-1 b) into hvec so that B_out
// survives back substitution, which overwrites that column with x.
static int snapshot_rhs(gauss_cuda_ctx* c)
{
    cudaStream_t s = c->s_main;
    for (int R = c->rank; R < c->nblk; R += c->world) {
        const int g0 = R * c->NB, l0 = blk_local_row(c, R), rows = blk_rows(c, R);
        launch_get_column(c->W + (size_t)l0 * c->ld, c->ld, c->np, c->hvec + g0, rows, s);
    }
    CK(cudaGetLastError());
    return GAUSS_CUDA_OK;
}

// ev[1] | computeGauss | ev[2] | snapshot of y | solveGauss | ev[3]
static int core_body(gauss_cuda_ctx* c)
{
    cudaStream_t s = c->s_main;
    int rc;
    CK(rec_event(c, c->ev[1], s));
    if ((rc = factor(c))) return rc;
    CK(rec_event(c, c->ev[2], s));
    if ((rc = snapshot_rhs(c))) return rc;
    if ((rc = backsub(c))) return rc;
    CK(rec_event(c, c->ev[3], s));
    c->factored = false;                     // the RHS column now holds x: solveGauss must not run a second time
    return GAUSS_CUDA_OK;
}

static int run_core(gauss_cuda_ctx* c)
{
    cudaStream_t s = c->s_main;
    const bool graphable = c->world == 1 && !c->pivot && !c->cp && c->use_graph && !c->graph_failed;
    if (graphable && !c->gexec) {
        const long long before = g_launch_count;
        c->capturing = true;
        cudaError_t e = cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
        int rc = e == cudaSuccess ? core_body(c) : GAUSS_CUDA_CUDA_ERR;
        cudaGraph_t graph = nullptr;
        if (e == cudaSuccess) e = cudaStreamEndCapture(s, &graph);
        c->capturing = false;
        if (rc == GAUSS_CUDA_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&c->gexec, graph, 0);
        if (graph) cudaGraphDestroy(graph);
        if (rc != GAUSS_CUDA_OK || e != cudaSuccess || !c->gexec) {
            c->gexec = nullptr;
            c->graph_failed = true;                      // fall back to plain launches for good
            cudaGetLastError();
        } else {
            c->graph_launches = g_launch_count - before;
            c->graph_upd_count = c->upd_count;
        }
        g_launch_count = before;
    }
    if (graphable && c->gexec) {
        CK(cudaMemsetAsync(c->status, 0, sizeof(int), s));   // (also inside the graph; harmless)
        c->upd_count = c->graph_upd_count;
        c->bcast_bytes = 0.0;
        CK(cudaGraphLaunch(c->gexec, s));
        g_launch_count += c->graph_launches;
        c->factored = false;                 // elimination and back substitution both ran (see core_body)
        c->last_graph = true;
        return GAUSS_CUDA_OK;
    }
    return core_body(c);
}

// Streams, events and the resident split-U buffers of the overlapped schedule, on first use.
static int stream_setup(gauss_cuda_ctx* c)
{
    if (c->s_copy) return GAUSS_CUDA_OK;
    CK(cudaStreamCreateWithFlags(&c->s_copy, cudaStreamNonBlocking));
    c->ev_chunk.assign((size_t)c->nblk, nullptr);
    for (auto& e : c->ev_chunk) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&c->ev_gate, cudaEventDisableTiming));
    CK(cudaEventCreate(&c->ev_h2d_done));
    if (c->mode == GAUSS_CUDA_MODE_3XTF32) {
        const size_t bytes = (size_t)c->nblk * c->ld * c->NB * sizeof(float);
        CK(cudaMalloc((void**)&c->ub_all_hi, bytes));
        CK(cudaMalloc((void**)&c->ub_all_lo, bytes));
        CK(cudaMemset(c->ub_all_hi, 0, bytes));
        CK(cudaMemset(c->ub_all_lo, 0, bytes));
    }
    return GAUSS_CUDA_OK;
}

extern "C" int gauss_cuda_solve(gauss_cuda_ctx* c, const float* A, const float* B, float* X, float* B_out, gauss_cuda_stats* st)
{
    if (!c || !A || !B || !X) return GAUSS_CUDA_BAD_ARG;
    if (!c->members.empty()) {
        const size_t n = (size_t)c->n;
        return group_call(c, [&](gauss_cuda_ctx* m, int r) {
            float* tx = c->group_tmp.data() + 2 * n * (size_t)r;
            return gauss_cuda_solve(m, A, B, r ? tx : X, B_out ? (r ? tx + n : B_out) : nullptr, r ? nullptr : st);
        });
    }
    CK(cudaSetDevice(c->device));
    g_launch_count = 0;
    cudaStream_t s = c->s_main;
    int rc;
    c->last_streamed = c->world == 1 && !c->pivot && !c->cp && c->stream_on && c->np >= c->stream_min_n && c->nblk >= 2;
    if (c->last_streamed && (rc = stream_setup(c))) return rc;
    CK(cudaEventRecord(c->ev[0], s));
    if (c->last_streamed) {
        CK(cudaEventRecord(c->ev[1], s));
        if ((rc = factor_streamed(c, A, B))) return rc;
        CK(cudaEventRecord(c->ev[2], s));
        if ((rc = snapshot_rhs(c))) return rc;
        if ((rc = backsub(c))) return rc;
        CK(cudaEventRecord(c->ev[3], s));
        c->factored = false;
    } else {
        if ((rc = upload_into(c, c->W, A, B))) return rc;
        if ((rc = run_core(c))) return rc;
    }
    if ((rc = download_results(c, X, B_out))) return rc;
    CK(cudaEventRecord(c->ev[4], s));
    rc = finish_stats(c, st, true, true, true, true);
    c->last_streamed = false;
    return rc;
}

extern "C" int gauss_cuda_upload(gauss_cuda_ctx* c, const float* A, const float* B)
{
    if (!c || !A || !B) return GAUSS_CUDA_BAD_ARG;
    if (!c->members.empty()) return group_call(c, [&](gauss_cuda_ctx* m, int) { return gauss_cuda_upload(m, A, B); });
    CK(cudaSetDevice(c->device));
    if (!c->W0) CK(cudaMalloc((void**)&c->W0, (size_t)std::max(c->rows_local, 1) * c->ld * sizeof(float)));
    int rc = upload_into(c, c->W0, A, B);
    if (rc) return rc;
    CK(cudaStreamSynchronize(c->s_main));
    c->resident = true;
    c->factored = false;
    return GAUSS_CUDA_OK;
}

// initMatrix (gauss_internal_input.c:50-60) generated on the device, straight into the resident copy: the fixture
// never exists on the host and "Application time" loses the 4 N^2-byte upload.  Works for every rank of a group.
extern "C" int gauss_cuda_init_matrix_device(gauss_cuda_ctx* c)
{
    if (!c) return GAUSS_CUDA_BAD_ARG;
    if (!c->members.empty()) return group_call(c, [&](gauss_cuda_ctx* m, int) { return gauss_cuda_init_matrix_device(m); });
    CK(cudaSetDevice(c->device));
    if (!c->W0) CK(cudaMalloc((void**)&c->W0, (size_t)std::max(c->rows_local, 1) * c->ld * sizeof(float)));
    cudaStream_t s = c->s_main;
    const int n = c->n;
    for (int R = c->rank; R < c->nblk; R += c->world) {
        const int g0 = R * c->NB, l0 = blk_local_row(c, R), rows = blk_rows(c, R);
        const int real = std::max(0, std::min(rows, n - g0));
        launch_init_fixture(c->W0, c->ld, l0, g0, real, n, c->np, s);
        if (real > 0) launch_pad_identity(c->W0, c->ld, l0, g0, real, n, n, s);
        if (rows > real) launch_pad_identity(c->W0, c->ld, l0 + real, g0 + real, rows - real, n, 0, s);
    }
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s));
    c->resident = true;
    c->factored = false;
    return GAUSS_CUDA_OK;
}

static int restore_working(gauss_cuda_ctx* c)
{
    CK(cudaMemcpyAsync(c->W, c->W0, (size_t)c->rows_local * c->ld * sizeof(float), cudaMemcpyDeviceToDevice, c->s_main));
    return GAUSS_CUDA_OK;
}

extern "C" int gauss_cuda_solve_resident(gauss_cuda_ctx* c, gauss_cuda_stats* st)
{
    if (!c) return GAUSS_CUDA_BAD_ARG;
    if (!c->members.empty()) return group_call(c, [&](gauss_cuda_ctx* m, int r) { return gauss_cuda_solve_resident(m, r ? nullptr : st); });
    if (!c->resident) return GAUSS_CUDA_NOT_READY;
    CK(cudaSetDevice(c->device));
    g_launch_count = 0;
    cudaStream_t s = c->s_main;
    int rc;
    CK(cudaEventRecord(c->ev[0], s));
    if ((rc = restore_working(c))) return rc;
    if ((rc = run_core(c))) return rc;
    CK(cudaEventRecord(c->ev[4], s));
    return finish_stats(c, st, true, true, true, false);
}

extern "C" int gauss_cuda_compute_gauss(gauss_cuda_ctx* c, gauss_cuda_stats* st)
{
    if (!c) return GAUSS_CUDA_BAD_ARG;
    if (!c->members.empty()) return group_call(c, [&](gauss_cuda_ctx* m, int r) { return gauss_cuda_compute_gauss(m, r ? nullptr : st); });
    if (!c->resident) return GAUSS_CUDA_NOT_READY;
    CK(cudaSetDevice(c->device));
    g_launch_count = 0;
    cudaStream_t s = c->s_main;
    int rc;
    CK(cudaEventRecord(c->ev[0], s));
    if ((rc = restore_working(c))) return rc;
    CK(cudaEventRecord(c->ev[1], s));
    if ((rc = factor(c))) return rc;
    CK(cudaEventRecord(c->ev[2], s));
    if ((rc = snapshot_rhs(c))) return rc;
    CK(cudaEventRecord(c->ev[3], s));
    CK(cudaEventRecord(c->ev[4], s));
    return finish_stats(c, st, true, true, false, false);
}

extern "C" int gauss_cuda_solve_gauss(gauss_cuda_ctx* c, gauss_cuda_stats* st)
{
    if (!c) return GAUSS_CUDA_BAD_ARG;
    if (!c->members.empty()) return group_call(c, [&](gauss_cuda_ctx* m, int r) { return gauss_cuda_solve_gauss(m, r ? nullptr : st); });
    if (!c->resident || !c->factored) return GAUSS_CUDA_NOT_READY;
    CK(cudaSetDevice(c->device));
    g_launch_count = 0;
    cudaStream_t s = c->s_main;
    int rc;
    CK(cudaEventRecord(c->ev[2], s));
    if ((rc = backsub(c))) return rc;
    CK(cudaEventRecord(c->ev[3], s));
    c->factored = false;                     // the RHS column now holds x: a second call would be wrong
    return finish_stats(c, st, false, false, true, false);
}

extern "C" int gauss_cuda_download(gauss_cuda_ctx* c, float* X, float* B_out)
{
    if (!c) return GAUSS_CUDA_BAD_ARG;
    if (!c->members.empty()) {
        const size_t n = (size_t)c->n;
        return group_call(c, [&](gauss_cuda_ctx* m, int r) {
            float* tx = c->group_tmp.data() + 2 * n * (size_t)r;
            return gauss_cuda_download(m, X ? (r ? tx : X) : nullptr, B_out ? (r ? tx + n : B_out) : nullptr);
        });
    }
    CK(cudaSetDevice(c->device));
    int rc = download_results(c, X, B_out);
    if (rc) return rc;
    CK(cudaStreamSynchronize(c->s_main));
    return GAUSS_CUDA_OK;
}

// Mixed-precision iterative refinement: the FP32 factors are the preconditioner, residuals and the solution
// are kept in double.  Each sweep: r = b - A x (double), d = U^-1 L^-1 (float)r with the stored factors, x += d.
extern "C" int gauss_cuda_solve_refined(gauss_cuda_ctx* c, const double* A, const double* B, double* X,
                                        int max_iters, double tol, gauss_cuda_refine_info* info)
{
    if (!c || !A || !B || !X || max_iters < 0) return GAUSS_CUDA_BAD_ARG;
    if (c->world != 1 || !c->members.empty()) {
        snprintf(g_err, sizeof g_err, "gauss_cuda_solve_refined: single-GPU contexts only");
        return GAUSS_CUDA_BAD_ARG;
    }
    if (c->pivot && (size_t)c->np * sizeof(float) > 200 * 1024) {
        snprintf(g_err, sizeof g_err, "gauss_cuda_solve_refined with pivoting: padded order %d exceeds 51200", c->np);
        return GAUSS_CUDA_BAD_ARG;
    }
    CK(cudaSetDevice(c->device));
    const int n = c->n, np = c->np, ld = c->ld;
    if (!c->A64) {
        CK(cudaMalloc((void**)&c->A64, (size_t)n * n * sizeof(double)));
        CK(cudaMalloc((void**)&c->v64, ((size_t)3 * n + 2) * sizeof(double)));
        CK(cudaMalloc((void**)&c->rf32, (size_t)3 * np * sizeof(float)));
    }
    double *b64 = c->v64, *x64 = c->v64 + n, *r64 = c->v64 + 2 * (size_t)n, *norms = c->v64 + 3 * (size_t)n;
    float *rf = c->rf32, *yf = c->rf32 + np, *df = c->rf32 + 2 * (size_t)np;
    g_launch_count = 0;
    cudaStream_t s = c->s_main;
    int rc;
    CK(cudaEventRecord(c->ev[0], s));
    CK(cudaMemcpyAsync(c->A64, A, (size_t)n * n * sizeof(double), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(b64, B, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    launch_f64_to_w(c->A64, b64, n, c->W, ld, np, s);
    launch_pad_identity(c->W, ld, 0, 0, n, n, n, s);
    if (np > n) launch_pad_identity(c->W, ld, n, n, np - n, n, 0, s);
    CK(cudaMemsetAsync(x64, 0, (size_t)n * sizeof(double), s));
    CK(cudaMemsetAsync(c->rf32, 0, (size_t)3 * np * sizeof(float), s));      // the padded tail of r stays zero
    if ((rc = run_core(c))) return rc;                                      // FP32 factors in W, x0 in xvec
    launch_refine_update(x64, c->xvec, n, norms, s);                        // x = (double)x0
    CK(cudaEventRecord(c->ev[4], s));
    rc = finish_stats(c, nullptr, true, true, true, false);
    if (rc) return rc;                                                      // zero pivot: nothing to refine
    if (c->pivot)                            // P A = L U with L in final row order: the earlier L columns get the later panels' swaps
        for (int k0 = kNb; k0 < np; k0 += kNb) launch_laswp_left(c->W, ld, k0, c->ipiv, s);
    double h[2] = {0.0, 0.0}, prev = 0.0;
    int it = 0, converged = 0;
    for (; it < max_iters; ++it) {
        launch_residual_f64(c->A64, x64, b64, r64, rf, n, s);
        if (c->pivot) launch_apply_pivots_vec(rf, c->ipiv, np, s);           // d = U^-1 L^-1 P r
        launch_trsv_factor_solve(c->W, ld, np, rf, yf, true, c->trsv_ctrl, s);
        launch_trsv_factor_solve(c->W, ld, np, yf, df, false, c->trsv_ctrl, s);
        launch_refine_update(x64, df, n, norms, s);
        CK(cudaMemcpyAsync(h, norms, sizeof h, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        const double rel = h[1] > 0.0 ? h[0] / h[1] : h[0];
        if (rel <= tol) { ++it; prev = rel; converged = 1; break; }
        if (!(rel == rel) || (it > 0 && rel > 0.5 * prev)) {
            // the corrections no longer halve: either the iteration does not contract (cond(A) 2^-24 too close to 1),
            // or -- far below FP32 accuracy -- it has reached what double residuals can resolve for this matrix
            // (the reference's own error on orsreg_1 is 2e-10): 2 = stagnated at working precision, 0 = failed
            ++it; converged = (rel == rel && rel <= 1e-7) ? 2 : 0; prev = rel; break;
        }
        prev = rel;
    }
    launch_residual_f64(c->A64, x64, b64, r64, rf, n, s);                   // the residual of the returned x, for info
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(X, x64, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
    CK(cudaEventRecord(c->ev[5], s));
    CK(cudaStreamSynchronize(s));
    c->factored = false;
    if (info) {
        std::vector<double> r((size_t)n);
        CK(cudaMemcpy(r.data(), r64, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost));
        double rmax = 0.0;
        for (double v : r) rmax = std::max(rmax, std::fabs(v));
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]);
        info->iterations = it;
        info->converged = converged;
        info->last_correction = prev;
        info->residual_inf = rmax;
        info->factor_s = c->last.h2d_s + c->last.factor_s + c->last.backsub_s;
        info->refine_s = (double)ms * 1e-3;
        info->kernel_launches = g_launch_count;
    }
    return GAUSS_CUDA_OK;
}

extern "C" int gauss_cuda_set_pivoting(gauss_cuda_ctx* c, int enabled)
{
    if (!c) return GAUSS_CUDA_BAD_ARG;
    if (c->world != 1) { snprintf(g_err, sizeof g_err, "partial pivoting is a single-GPU path"); return GAUSS_CUDA_BAD_ARG; }
    CK(cudaSetDevice(c->device));
    if (enabled && !c->pt) {
        CK(cudaMalloc((void**)&c->pt, (size_t)kNb * c->np * sizeof(float)));
        CK(cudaMalloc((void**)&c->ipiv, (size_t)c->np * sizeof(int)));
        CK(cudaMemset(c->ipiv, 0, (size_t)c->np * sizeof(int)));
    }
    c->pivot = enabled != 0;
    c->factored = false;
    return GAUSS_CUDA_OK;
}

extern "C" int gauss_cuda_get_pivots(gauss_cuda_ctx* c, int* pivot_rows)
{
    if (!c || !pivot_rows) return GAUSS_CUDA_BAD_ARG;
    if (!c->ipiv) return GAUSS_CUDA_NOT_READY;
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->s_main));
    CK(cudaMemcpy(pivot_rows, c->ipiv, (size_t)c->n * sizeof(int), cudaMemcpyDeviceToHost));
    return GAUSS_CUDA_OK;
}

extern "C" int gauss_cuda_download_factors(gauss_cuda_ctx* c, float* LU)
{
    if (!c || !LU) return GAUSS_CUDA_BAD_ARG;
    if (c->world != 1) return GAUSS_CUDA_BAD_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpy2DAsync(LU, (size_t)c->n * sizeof(float), c->W, (size_t)c->ld * sizeof(float),
                         (size_t)c->n * sizeof(float), c->n, cudaMemcpyDeviceToHost, c->s_main));
    CK(cudaStreamSynchronize(c->s_main));
    return GAUSS_CUDA_OK;
}

// ------------------------------------------------------------------------------------------
// FP32 matrix multiply
// ------------------------------------------------------------------------------------------
struct MatmulScratch {            // per device, grown on demand, kept until process exit
    float* buf = nullptr;
    size_t bytes = 0;
    // every call on a device shares `buf`; calls on different streams (or from different host threads) are put in
    // order with this event -- the next call's stream waits for the previous call's kernels -- so they can never
    // overwrite each other's operand copies (the reference's gpu_matmul<<<>>> launch has no such restriction either)
    cudaEvent_t last_done = nullptr;
    cudaStream_t last_stream = nullptr;
    bool used = false;
};
static MatmulScratch g_mm[16];
static std::mutex g_mm_mutex;

static int mm_scratch(int dev, size_t bytes, float** out)
{
    MatmulScratch& sc = g_mm[dev & 15];
    if (sc.bytes < bytes) {
        if (sc.buf) { CK(cudaDeviceSynchronize()); CK(cudaFree(sc.buf)); sc.buf = nullptr; sc.bytes = 0; }
        CK(cudaMalloc((void**)&sc.buf, bytes));
        sc.bytes = bytes;
    }
    *out = sc.buf;
    return GAUSS_CUDA_OK;
}

// Core of both entry points: C = A*B on device pointers with caller-provided scratch (3xTF32 mode:
// 4 * n * kp floats, kp = n rounded up to 32; FFMA mode with n % 16 != 0: 3 padded copies).
size_t matmul_scratch_bytes(int n, int mode)
{
    if (mode == GAUSS_CUDA_MODE_3XTF32) {
        const size_t kp = (size_t)round_up(n, 32);
        size_t b = 4 * (size_t)n * kp * sizeof(float);
        if (n % 4) b += (size_t)n * round_up(n, 4) * sizeof(float);     // padded C
        return b;
    }
    if (n % 16 == 0) return 0;
    const size_t p = (size_t)round_up(n, 16);
    return 3 * p * p * sizeof(float);
}

int matmul_on_device(const float* dA, const float* dB, float* dC, int n, int mode, float* scratch, cudaStream_t s)
{
    if (mode == GAUSS_CUDA_MODE_FFMA) {
        if (n % 16 == 0) {
            launch_sgemm_nn(dA, n, dB, n, dC, n, n, n, n, 1.f, 0.f, s);
        } else {
            const int p = round_up(n, 16);
            float* pa = scratch, *pb = scratch + (size_t)p * p, *pc = pb + (size_t)p * p;
            CK(cudaMemsetAsync(pa, 0, 2 * (size_t)p * p * sizeof(float), s));
            CK(cudaMemcpy2DAsync(pa, (size_t)p * 4, dA, (size_t)n * 4, (size_t)n * 4, n, cudaMemcpyDeviceToDevice, s));
            CK(cudaMemcpy2DAsync(pb, (size_t)p * 4, dB, (size_t)n * 4, (size_t)n * 4, n, cudaMemcpyDeviceToDevice, s));
            launch_sgemm_nn(pa, p, pb, p, pc, p, n, p, p, 1.f, 0.f, s);
            CK(cudaMemcpy2DAsync(dC, (size_t)n * 4, pc, (size_t)p * 4, (size_t)n * 4, n, cudaMemcpyDeviceToDevice, s));
        }
        CK(cudaGetLastError());
        return GAUSS_CUDA_OK;
    }
    const int kp = round_up(n, 32);
    float* ahi = scratch, *alo = ahi + (size_t)n * kp, *bhi = alo + (size_t)n * kp, *blo = bhi + (size_t)n * kp;
    launch_split_rowmajor(dA, n, ahi, alo, kp, n, n, s);
    launch_split_transpose(dB, n, bhi, blo, kp, n, n, kp, s);
    float* cdst = dC;
    int ldc = n;
    if (n % 4) { cdst = blo + (size_t)n * kp; ldc = round_up(n, 4); }
    int rc = tc_gemm_nt(ahi, alo, kp, bhi, blo, kp, cdst, ldc, n, n, kp, +1, 0, s);
    if (rc) return cuda_fail((cudaError_t)rc, "tc_gemm_nt", __LINE__);
    if (cdst != dC) CK(cudaMemcpy2DAsync(dC, (size_t)n * 4, cdst, (size_t)ldc * 4, (size_t)n * 4, n, cudaMemcpyDeviceToDevice, s));
    CK(cudaGetLastError());
    return GAUSS_CUDA_OK;
}

extern "C" int matmul_cuda_device(const float* dA, const float* dB, float* dC, int n, int mode, void* stream)
{
    if (!dA || !dB || !dC || n < 1) return GAUSS_CUDA_BAD_ARG;
    if (mode != GAUSS_CUDA_MODE_3XTF32 && mode != GAUSS_CUDA_MODE_FFMA) return GAUSS_CUDA_BAD_ARG;
    if (gauss_cuda_device_count() < 1) { snprintf(g_err, sizeof g_err, "no CUDA device (no CPU fallback)"); return GAUSS_CUDA_CUDA_ERR; }
    int dev = 0;
    CK(cudaGetDevice(&dev));
    simt_setup_device();
    if (mode == GAUSS_CUDA_MODE_3XTF32) {
        if (!tc_available()) { snprintf(g_err, sizeof g_err, "GAUSS_CUDA_MODE_3XTF32 needs an sm_100 device"); return GAUSS_CUDA_CUDA_ERR; }
        int rc = tc_setup_device();
        if (rc) return rc;
    }
    float* scratch = nullptr;
    const size_t need = matmul_scratch_bytes(n, mode);
    if (!need) return matmul_on_device(dA, dB, dC, n, mode, nullptr, (cudaStream_t)stream);
    std::lock_guard<std::mutex> lock(g_mm_mutex);
    MatmulScratch& sc = g_mm[dev & 15];
    { int rc = mm_scratch(dev, need, &scratch); if (rc) return rc; }
    if (!sc.last_done) CK(cudaEventCreateWithFlags(&sc.last_done, cudaEventDisableTiming));
    if (sc.used && sc.last_stream != (cudaStream_t)stream) CK(cudaStreamWaitEvent((cudaStream_t)stream, sc.last_done, 0));
    int rc = matmul_on_device(dA, dB, dC, n, mode, scratch, (cudaStream_t)stream);
    if (rc) return rc;
    CK(cudaEventRecord(sc.last_done, (cudaStream_t)stream));
    sc.last_stream = (cudaStream_t)stream;
    sc.used = true;
    return GAUSS_CUDA_OK;
}

extern "C" int matmul_cuda(const float* A, const float* B, float* C, int n, int mode, double* kernel_s)
{
    if (!A || !B || !C || n < 1) return GAUSS_CUDA_BAD_ARG;
    if (gauss_cuda_device_count() < 1) { snprintf(g_err, sizeof g_err, "no CUDA device (no CPU fallback)"); return GAUSS_CUDA_CUDA_ERR; }
    g_launch_count = 0;
    const size_t bytes = (size_t)n * n * sizeof(float);
    float* d = nullptr;
    CK(cudaMalloc((void**)&d, 3 * bytes));
    float* dA = d, *dB = d + (size_t)n * n, *dC = dB + (size_t)n * n;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int rc = GAUSS_CUDA_OK;
    cudaError_t e = cudaMemcpy(dA, A, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dB, B, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        cudaEventRecord(e0, 0);
        rc = matmul_cuda_device(dA, dB, dC, n, mode, nullptr);
        cudaEventRecord(e1, 0);
        if (rc == GAUSS_CUDA_OK) e = cudaMemcpy(C, dC, bytes, cudaMemcpyDeviceToHost);
    }
    if (rc == GAUSS_CUDA_OK && e == cudaSuccess && kernel_s) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        *kernel_s = (double)ms * 1e-3;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d);
    if (rc) return rc;
    if (e != cudaSuccess) return cuda_fail(e, "matmul_cuda copies", __LINE__);
    return GAUSS_CUDA_OK;
}

// ------------------------------------------------------------------------------------------
// matmul for callers that may not allocate
// ------------------------------------------------------------------------------------------
// Some hosts cannot let a library call cudaMalloc: an application that sized its device memory at
// start-up and forbids driver allocations afterwards (a late cudaMalloc can fail under memory
// pressure and it synchronises the whole device), or a latency-bound loop that must not block.
// For them the scratch of the multiply lives in the array below instead of a cudaMalloc'd buffer.
// It is part of the library image: EVERY GPU that loads libgauss_cuda.so carries these 20 MiB
// until the process exits, whether or not the entry point is ever called.
constexpr int    kNoAllocMaxN = 1024;
constexpr size_t kNoAllocFloats = 5u * 1024u * 1024u;               // 4 split operand copies + one padded C, n <= 1024
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocFloats];

// In host code `g_noalloc_scratch` is only a handle; the device address has to be asked for, once per
// device, with that device current.  Cached here after the first successful query.
static float* g_noalloc_addr[16] = {nullptr};

extern "C" int matmul_cuda_device_noalloc(const float* dA, const float* dB, float* dC, int n, int mode, void* stream)
{
    if (!dA || !dB || !dC || n < 1) return GAUSS_CUDA_BAD_ARG;
    if (mode != GAUSS_CUDA_MODE_3XTF32 && mode != GAUSS_CUDA_MODE_FFMA) return GAUSS_CUDA_BAD_ARG;
    if (n > kNoAllocMaxN) {
        snprintf(g_err, sizeof g_err, "matmul_cuda_device_noalloc: n = %d exceeds the built-in scratch (n <= %d); "
                 "use matmul_cuda_device, which allocates", n, kNoAllocMaxN);
        return GAUSS_CUDA_BAD_ARG;
    }
    if (gauss_cuda_device_count() < 1) { snprintf(g_err, sizeof g_err, "no CUDA device (no CPU fallback)"); return GAUSS_CUDA_CUDA_ERR; }
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 16) {
        snprintf(g_err, sizeof g_err, "matmul_cuda_device_noalloc: device ordinal %d is outside the 16 this library keeps set-up state for", dev);
        return GAUSS_CUDA_BAD_ARG;
    }
    if (matmul_scratch_bytes(n, mode) > kNoAllocFloats * sizeof(float)) {   // cannot happen for n <= 1024
        snprintf(g_err, sizeof g_err, "matmul_cuda_device_noalloc: scratch for n = %d does not fit the built-in array", n);
        return GAUSS_CUDA_BAD_ARG;
    }
    if (mode == GAUSS_CUDA_MODE_3XTF32 && !tc_available()) {
        snprintf(g_err, sizeof g_err, "GAUSS_CUDA_MODE_3XTF32 needs an sm_100 device");
        return GAUSS_CUDA_CUDA_ERR;
    }
    if (!g_noalloc_addr[dev]) {
        // first call on this device: one-time set-up (symbol address, kernel attributes, TMA encoder).  The tile
        // engine's process-wide switches are left as they are when another caller has already set them up.
        float* p = nullptr;
        CK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        simt_setup_device();
        if (tc_available()) {
            int rc = tc_setup_device(/*reread_env=*/false);
            if (rc) return rc;
        }
        g_noalloc_addr[dev] = p;
    }
    // Every region of the scratch that this call reads is written by this call first (the operand
    // splits include their zero K-padding; the padded FFMA path clears its operand copies), so the
    // previous call's contents never leak into the result.
    return matmul_on_device(dA, dB, dC, n, mode, g_noalloc_addr[dev], (cudaStream_t)stream);
}
