// Internal declarations shared by the translation units of libgauss_cuda.so.
// Nothing here is part of the C ABI (include/gauss_cuda.h is).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gb200 {

constexpr int kNb = 128;          // inner block: diagonal factor / substitution kernels
constexpr int kPadCols = 32;      // zero columns after the RHS column (keeps rows 128 B aligned)

// ---- SIMT kernels (lu_simt.cu) ---------------------------------------------------------

// In-place Crout factor of the 128x128 block at `blk` (leading dimension ld):
// U unit-upper above the diagonal, L (pivots included) on and below it.
// `status` receives (global_row + 1) of the first zero / non-finite pivot (0 = none).
// U-row solve of (a12, ncols) and L-col solve of (a21, nrows) against the same factored diagonal block d11, in one
// launch where the tensor-core panel kernels are in use (otherwise two launches)
void launch_urow_lcol(const float* d11, int ldd,
                      float* a12, int ld12, int ncols, float* ub_hi, float* ub_lo, int ldub, int k_ub,
                      float* a21, int ld21, int nrows, float* la_hi, float* la_lo, int ldla, int k_la, cudaStream_t s);
void launch_diag_factor(float* blk, int ld, int global_row0, int* status, cudaStream_t s);

// U-row: X = L11^-1 * A12 for the 128-row block `rows` over `ncols` columns starting at
// `a12` (ld); L11 is the lower triangle (with pivots) of the 128x128 block at `l11` (ldl).
// If ub_hi != nullptr also writes the TF32 hi/lo split of X TRANSPOSED:
// ub_{hi,lo}[(col0_ub + c) * ldub + k_ub + r] for column c, row r (B-role operand, K-major).
void launch_urow(const float* l11, int ldl, float* a12, int ld, int ncols,
                 float* ub_hi, float* ub_lo, int ldub, int k_ub, cudaStream_t s);

// L-col: X = A21 * U11^-1 for `nrows` rows starting at `a21` (ld), 128 columns; U11 is the
// strict upper triangle (unit diagonal implied) of the block at `u11` (ldu).
// If la_hi != nullptr also writes the hi/lo split of X: la_{hi,lo}[r * ldla + k_la + c].
void launch_lcol(const float* u11, int ldu, float* a21, int ld, int nrows,
                 float* la_hi, float* la_lo, int ldla, int k_la, cudaStream_t s);

// FP32 FFMA GEMM: C[m x n] = beta*C + alpha*A[m x k]*B[k x n], row-major, alpha in {+1,-1}, beta in {0,1}.
// m, n arbitrary (n and all leading dimensions multiples of 4, pointers 16 B aligned), k multiple of 16.
void launch_sgemm_nn(const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                     int m, int n, int k, float alpha, float beta, cudaStream_t s);

// Whole back substitution of an [nrows x nrows] unit-upper system in one launch: x = U^-1 * W[:, ycol]
// (W points at the system's element (0,0); ycol is relative to it).  ctrl: nrows/128 + 1 ints of device
// scratch (ticket + per-block flags), reset here.  W is not modified.
void launch_trsv_fused(const float* W, int ld, int nrows, int ycol, float* x, int* ctrl, cudaStream_t s);
void launch_trsv_gemv_wide(float* u, int ld, int ycol, const float* x, int nrows, int kw, cudaStream_t s);
// Solves with the stored factors for a new right-hand side (device vectors of nrows floats): y = L^-1 rhs
// (lower, L carries the pivots) or x = U^-1 rhs (unit upper).  W is not modified.
void launch_trsv_factor_solve(const float* W, int ld, int nrows, const float* rhs, float* out, bool lower, int* ctrl, cudaStream_t s);
// FP64 helpers of the iterative-refinement entry point (see gauss_cuda_solve_refined)
void launch_f64_to_w(const double* A, const double* b, int n, float* w, int ld, int ycol, cudaStream_t s);
void launch_residual_f64(const double* A, const double* x, const double* b, double* r, float* rf, int n, cudaStream_t s);
void launch_refine_update(double* x, const float* d, int n, double* out, cudaStream_t s);

// Layout helpers: scatter the dense RHS into column `col` / gather it back; identity padding.
void launch_init_fixture(float* w, int ld, int row0_local, int g0, int nrows, int n, int ycol, cudaStream_t s);
void launch_set_column(float* w, int ld, int col, const float* src, int nrows, cudaStream_t s);
void launch_get_column(const float* w, int ld, int col, float* dst, int nrows, cudaStream_t s);
void launch_pad_identity(float* w, int ld, int row0_local, int row0_global, int nrows, int n, int col0, cudaStream_t s);

// matmul operand preparation for the tensor-core path: A -> hi/lo (row-major), B -> hi/lo transposed.
void launch_split_rowmajor(const float* src, int lds, float* hi, float* lo, int ldd, int rows, int cols, cudaStream_t s);
// split_transpose: hi/lo[c*ldd + r] = split(src[r*lds + c]) for c < cols, r < rext (source rows >= rows read as 0).
void launch_split_transpose(const float* src, int lds, float* hi, float* lo, int ldd, int rows, int cols, int rext, cudaStream_t s);
void simt_setup_device();     // per-device dynamic shared memory opt-ins of the SIMT kernels
void panel_setup_device();    // same for lu_panel.cu (called by simt_setup_device)
int  tc_setup_device(bool reread_env = true);   // same for the tile engine; returns a GAUSS_CUDA_* status.  reread_env = false leaves the
                                                // process-wide GAUSS_CUDA_TC_* switches alone once some caller has read them

// ---- critical-path schedule helpers (lu_cp.cu) ------------------------------------------------
// E[0:2kb, 0:2kb] <- [[A11, I], [I, 0]]; after the in-panel chain on E: [[L11|U11, inv(L11)], [inv(U11), *]].
void launch_ext_prepare(const float* a11, int ld, float* E, int lde, int kb, cudaStream_t s);
// L11|U11 back into the matrix; il_* = hi/lo split of inv(L11) (row-major, ldo); iut_* = hi/lo split of inv(U11)^T.
// peer_* (may be null): the next owner's copy of iut_* in peer-mapped memory; peer_flag (may be null): word in the peer's
// memory set to `seq` once every CTA's stores are out; counter: one zeroed device word per launch site.
void launch_ext_finish(const float* E, int lde, int kb, float* a11, int ld,
                       float* il_hi, float* il_lo, float* iut_hi, float* iut_lo, int ldo,
                       float* peer_hi, float* peer_lo, unsigned* counter, unsigned* peer_flag, unsigned seq, float* iu_raw, cudaStream_t s);
// x_K = inv(U11) y_K on the owner, stored locally and into every peer's receive vector, then the peers' flags
struct BsPeers { float* x[8]; unsigned* flag[8]; int n; };
void launch_bs_solve_push(const float* iu, int ldiu, const float* ycol, int ld, int kb, float* x_local, const BsPeers& peers,
                          unsigned* counter, unsigned seq, cudaStream_t s);
// hi/lo split of a row-major block into local (and peer) operand buffers, then the peer's flag
void launch_crit_split(const float* src, int lds, float* hi, float* lo, int ldd, int rows, int cols,
                       float* peer_hi, float* peer_lo, unsigned* counter, unsigned* peer_flag, unsigned seq, cudaStream_t s);
void launch_wait_flag(const unsigned* flag, unsigned value, cudaStream_t s);

// ---- partial-pivoting panel (lu_pivot.cu) --------------------------------------------------
// Factors the panel W[k0.., k0..k0+128) of an [np x ld] single-GPU system with row interchanges
// (gauss_external_input.c:123-148, 161-181); see lu_pivot.cu.  la_* (may be null): split of the rows
// below the diagonal block, la[(i-128)*ldla + c].
void launch_laswp_left(float* W, int ld, int k0, const int* ipiv, cudaStream_t s);      // panel k0's swaps on the columns left of it
int  launch_apply_pivots_vec(float* v, const int* ipiv, int np, cudaStream_t s);       // v <- P v; non-zero: np too large (> 51200)
void launch_panel_pivot(float* W, int ld, int np, int k0, float* pt, int mp, int* ipiv, int* status,
                        float* la_hi, float* la_lo, int ldla, cudaStream_t s);

// ---- tensor-core tile engine (tc_gemm.cu) -----------------------------------------------

// C[m x n] (ldc) = beta*C - or + (Ahi+Alo)[m x k] * (Bhi+Blo)[n x k]^T as three TF32 MMAs per k-slice.
// A* are [m x k] K-major (lda), B* are [n x k] K-major (ldb); sign = -1 subtracts, +1 adds; beta in {0,1}.
// k multiple of 32.  Returns a cudaError_t-compatible code (0 = ok).
int tc_gemm_nt(const float* Ahi, const float* Alo, int lda, const float* Bhi, const float* Blo, int ldb,
               float* C, int ldc, int m, int n, int k, int sign, int beta, cudaStream_t s);
bool tc_available();
// Scheduler slots (2 * 64 zeroed ints of device memory) used by this thread's following tc_gemm_nt calls;
// nullptr = the library's own ring.
void tc_set_sched_slots(int* slots);
void tc_set_pair_ok(bool ok);       // false: this thread's launches must use the dynamic-scheduler kernel (multi-GPU)
// > 0: CTAs of this thread's following tc_gemm_nt launches exit after that many tiles (the grid grows to match),
// so kernels of other streams get SMs while a long launch runs.  0 = persistent CTAs.
void tc_set_max_tiles(int max_tiles);
// > 0: this thread's following tc_gemm_nt launches use at most (SM count - n) SMs, so that the small kernels of a
// latency-bound chain on another stream always find a free SM while a long update runs.
void tc_set_reserve_sms(int n);

extern thread_local long long g_launch_count;   // kernels launched by this thread since last reset

// ---- programmatic dependent launch (PDL) ----------------------------------------------------
// The panel chain is a string of short dependent kernels; between two of them the GPU otherwise pays the drain of the
// first, the launch of the second and the second's prologue (barrier / TMEM set-up, tensor-map fetch).  Kernels of the
// chain therefore (a) release their dependents at once (griddepcontrol.launch_dependents at the top) and (b) wait for
// their predecessors' memory only where they first touch global data (griddepcontrol.wait).  Both instructions are
// no-ops in a launch without the attribute, which is how every other launch of the same kernels runs.
extern thread_local bool t_pdl;                 // this thread's following launch_k calls carry the PDL attribute
struct PdlScope { bool prev; explicit PdlScope(bool on) : prev(t_pdl) { t_pdl = on; } ~PdlScope() { t_pdl = prev; } };
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = t_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

} // namespace gb200
