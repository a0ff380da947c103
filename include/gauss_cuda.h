/*
 * gauss_cuda.h -- C ABI of the B200-native dense Gaussian-elimination solver and
 * FP32 matrix multiply (libgauss_cuda.so).  Plain C: pointers, sizes, status codes.
 *
 * The reference (svdeepak99/Gaussian_Elimination-CUDA-OpenMP-MPI-Pthreads) has no
 * library interface: its boundary is a source-level convention -- file-scope
 * globals plus void functions taking `int nsize`
 *   (OpenMP_and_MPI/gauss_openmp/gauss_internal_input.c:15-16, :20-43 allocate_memory,
 *    :50-60 initMatrix, :66-112 getPivot, :118-138 computeGauss, :144-159 solveGauss)
 * and, for the matrix multiply, one kernel launch on device pointers
 *   (CUDA_and_OpenMP/Version-2/cuda_matmul.cu:89-101 gpu_matmul, :140-160 host sequence).
 * Each entry point below names the reference interface it replaces.
 *
 * Conventions: every function returns a GAUSS_CUDA_* status and never calls exit();
 * host buffers are caller-owned; a context owns everything on the device and is not
 * thread-safe; NULL contexts are rejected (free(NULL) is a no-op).  There is NO CPU
 * fallback: without a CUDA device every compute call returns GAUSS_CUDA_CUDA_ERR.
 *
 * Semantics kept from the reference: no pivoting, the pivot row is normalised so U
 * has a unit diagonal and L carries the pivots (Crout), B after elimination is
 * L^-1 b, back substitution has no divisions.  One narrowing: the reference swaps
 * rows when a pivot is EXACTLY 0.0 (gauss_internal_input.c:75-97); this solver
 * reports GAUSS_CUDA_ZERO_PIVOT instead, which the harness maps to the reference's
 * "The matrix is singular" + exit(-1) (gauss_internal_input.c:86-89).
 */
#ifndef GAUSS_CUDA_H
#define GAUSS_CUDA_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gauss_cuda_ctx gauss_cuda_ctx;   /* opaque: device buffers, streams, NCCL communicator, peer mappings;
                                                 * for ngpus > 1 a handle on one such context per device */

enum {
    GAUSS_CUDA_OK         = 0,
    GAUSS_CUDA_ZERO_PIVOT = 1,   /* a pivot was 0 or non-finite; stats.zero_pivot_row says where */
    GAUSS_CUDA_BAD_ARG    = 2,
    GAUSS_CUDA_CUDA_ERR   = 3,   /* see gauss_cuda_last_error() */
    GAUSS_CUDA_NCCL_ERR   = 4,
    GAUSS_CUDA_NOT_READY  = 5    /* device-resident call before gauss_cuda_upload */
};

enum {
    GAUSS_CUDA_MODE_3XTF32 = 0,  /* trailing update / matmul on tcgen05 tensor cores, split 3xTF32 */
    GAUSS_CUDA_MODE_FFMA   = 1   /* same algorithm, pure FP32 FFMA kernels (comparison variant) */
};

typedef struct {
    double h2d_s, factor_s, backsub_s, d2h_s;   /* CUDA-event times of the last call, seconds */
    double bcast_bytes;                         /* bytes this rank received from the other ranks in the last call (NCCL
                                                 * broadcasts of the panels plus the x_K written into its memory by peers) */
    int    zero_pivot_row;                      /* -1, or the global row of the first bad pivot */
    int    n, n_padded, nb_outer, nb_inner, mode, rank, world;
    long long kernel_launches;                  /* kernels this library launched in the last call */
    /* the dominant kernel (trailing update A22 -= L21*U12, one launch per outer block), timed with CUDA
     * events on the library's stream: summed duration, algorithmic flops (2*m*n*k per launch), launches */
    double update_s, update_flops;
    int    update_launches;
    /* 1: gauss_cuda_solve ran the overlapped schedule (rows uploaded in chunks beside the elimination; single
     * GPU, n >= 16384 unless GAUSS_CUDA_STREAM / GAUSS_CUDA_STREAM_MIN_N / GAUSS_CUDA_STREAM_CHUNKS say
     * otherwise).  h2d_s is then the time until the last chunk had arrived and factor_s (which contains the
     * waiting for chunks) overlaps it; results are bit-identical to the upload-first schedule. */
    int    overlapped_upload;
    /* 1: the elimination + back substitution of this call ran as one replay of the context's captured CUDA graph
     * (single GPU, no pivoting); 0: plain stream launches (multi-GPU, pivoting, the overlapped upload, GAUSS_CUDA_GRAPH=0,
     * or graph capture failed on this context -- in which case every later call of the context stays on plain launches) */
    int    graph_replayed;
} gauss_cuda_stats;

/* ---- Gaussian elimination ------------------------------------------------------------- */

/* Replaces allocate_memory(size) (gauss_internal_input.c:20-43) for the device side.
 * n: matrix order (any n >= 1); ngpus: GPUs driven by THIS process: 1, or G > 1 for devices 0..G-1 of this node
 * (the reference's multi-worker run is one program as well: mpirun ... gauss_internal_input, OpenMP_and_MPI/README.txt:23;
 * rows are owned block-cyclically as in gauss_cuda_init_dist, each device is driven by a host thread of the library
 * during a call, the devices exchange panels over NVLink through NCCL and direct peer access; the context is used
 * exactly like a single-GPU one by the calling thread; pivoting, refinement and factor download stay single-GPU);
 * nb: outer block, 0 = auto, else a multiple of 128; mode: GAUSS_CUDA_MODE_*. */
int gauss_cuda_init(gauss_cuda_ctx **ctx, int n, int ngpus, int nb, int mode);

/* Multi-GPU form, one process per GPU: the reference's MPI row distribution
 * (OpenMP_and_MPI/gauss_mpi/gauss_internal_input.c:140-155) becomes block-cyclic row
 * panels (row block r lives on rank r mod world) and its per-step
 * MPI_Bcast(&matrix[i][i+1], ...) (gauss_mpi/gauss_internal_input.c:137) becomes one NCCL
 * broadcast of the factored row panel per outer block.  `nccl_id` is 128 bytes obtained
 * from gauss_cuda_unique_id() on rank 0 and handed to every rank by the launcher
 * (torch.distributed, MPI, a file ...).  Collective: every rank must call it.  In 3xTF32 mode the ranks also map each
 * other's receive buffers (CUDA IPC; needs peer access between the GPUs, as NVLink / NVSwitch give): what the next
 * panel needs first and the solution blocks of back substitution are stored there by the producing kernels themselves;
 * without peer access the same data travels by NCCL only. */
int gauss_cuda_unique_id(void *nccl_id_128);
int gauss_cuda_init_dist(gauss_cuda_ctx **ctx, int n, int rank, int world, int device,
                         const void *nccl_id_128, int nb, int mode);

/* gauss() of the north_star = initMatrix's data in, computeGauss + solveGauss
 * (gauss_internal_input.c:118-138, :144-159) on the GPU, X out.
 *   A      host, row-major n*n, ld = n, not modified (what matrix[0] holds, :30-33);
 *          in the multi-GPU form every rank passes the same-shaped buffer and only the
 *          rows it owns are read
 *   B      host, n                       (the reference's B, :57)
 *   X      host, n, out                  (the reference's C, :156-158)
 *   B_out  host, n, out, may be NULL     (B after elimination, printed by VERIFY, :225-228)
 *   st     may be NULL
 * Collective in the multi-GPU form. */
int gauss_cuda_solve(gauss_cuda_ctx *ctx, const float *A, const float *B,
                     float *X, float *B_out, gauss_cuda_stats *st);

/* Device-resident form, for callers that keep the system in HBM between solves
 * (and for kernel-only timing): upload once, then each gauss_cuda_solve_resident
 * restores the working matrix from the resident copy (device-to-device), eliminates
 * and back-substitutes entirely on the device; download fetches X / B_out. */
int gauss_cuda_upload(gauss_cuda_ctx *ctx, const float *A, const float *B);
int gauss_cuda_solve_resident(gauss_cuda_ctx *ctx, gauss_cuda_stats *st);
int gauss_cuda_download(gauss_cuda_ctx *ctx, float *X, float *B_out);

/* The two halves separately, as the reference exposes them: computeGauss
 * (gauss_internal_input.c:118-138) then solveGauss (:144-159), on the resident system. */
/* initMatrix(nsize) (gauss_internal_input.c:50-60: matrix[i][j] = 2 (min(i,j)+1), B[i] = i) generated on the device,
 * in place of gauss_host_init_matrix + gauss_cuda_upload: the fixture never exists on the host, so the reference's
 * timed window (init + elimination, :211-217) carries no 4 n^2-byte upload.  Every rank of a group fills its own rows. */
int gauss_cuda_init_matrix_device(gauss_cuda_ctx *ctx);
int gauss_cuda_compute_gauss(gauss_cuda_ctx *ctx, gauss_cuda_stats *st);
int gauss_cuda_solve_gauss(gauss_cuda_ctx *ctx, gauss_cuda_stats *st);

/* Reads back the factored matrix (unit-upper U above the diagonal, L with the pivots on
 * and below it -- the reference zeroes L, gauss_internal_input.c:131, nothing reads it).
 * LU: host, row-major n*n.  Single-GPU contexts only. */
int gauss_cuda_download_factors(gauss_cuda_ctx *ctx, float *LU);

/* Partial pivoting, the reference's EXTERNAL-input variant (gauss_openmp/gauss_external_input.c:123-148
 * getPivot: the row with the largest |a[i][cur]|, i >= cur, is swapped in; :161-169 the pivot row is
 * divided by the pivot).  enabled != 0 switches every later solve of this context to it; single-GPU
 * contexts only (GAUSS_CUDA_BAD_ARG otherwise).  A zero column is reported as GAUSS_CUDA_ZERO_PIVOT
 * (the reference prints "The matrix is singular", :133-136).  gauss_cuda_get_pivots returns, for every
 * pivot step i < n, the row that was swapped into position i during the last elimination. */
int gauss_cuda_set_pivoting(gauss_cuda_ctx *ctx, int enabled);
int gauss_cuda_get_pivots(gauss_cuda_ctx *ctx, int *pivot_rows /* host, n */);

/* Double-precision results from the FP32 factors.  The reference computes in double
 * (double **matrix, *B, *C, gauss_internal_input.c:15-16); this entry takes and returns doubles and reaches the
 * reference's accuracy by mixed-precision iterative refinement: A is rounded to FP32 and eliminated exactly as in
 * gauss_cuda_solve (computeGauss :118-141), then x is corrected in sweeps  r = b - A x  (double, on the device),
 * d = U^-1 L^-1 r  (the stored FP32 factors, two triangular solves -- solveGauss :144-159 and its forward
 * counterpart), x += d, until  max|d| <= tol * max|x|,  the corrections stop halving, or max_iters sweeps ran.
 * Converges when cond(A) * 2^-24 is well below 1 (the reference's own inputs are); `converged` says whether the
 * tolerance was met.  A, B, X: host, n*n / n / n doubles.  Single-GPU contexts; with partial pivoting switched on
 * (gauss_cuda_set_pivoting) the sweeps solve with P A = L U -- the external-input program in the reference's own
 * precision (gauss_external_input.c prints Error: 4.8e-15 for jpwh_991) -- for padded orders up to 51200.
 * Device memory for the double copy of A (8 n^2 bytes) is allocated on the first call and kept by the context. */
typedef struct gauss_cuda_refine_info {
    int iterations;            /* refinement sweeps done */
    int converged;             /* 1: last max|d|/max|x| <= tol; 2: the corrections stopped halving below 1e-7, i.e. the
                                * answer is as good as double residuals can make it for this matrix; 0: neither */
    double last_correction;    /* max|d| / max|x| of the last sweep */
    double residual_inf;       /* max_i |b - A x|_i of the returned x, in double */
    double factor_s;           /* upload + FP32 elimination + first back substitution, seconds */
    double refine_s;           /* all sweeps + download, seconds */
    long long kernel_launches;
} gauss_cuda_refine_info;
int gauss_cuda_solve_refined(gauss_cuda_ctx *ctx, const double *A, const double *B, double *X,
                             int max_iters, double tol, gauss_cuda_refine_info *info /* may be NULL */);

void gauss_cuda_free(gauss_cuda_ctx *ctx);

/* ---- FP32 matrix multiply -------------------------------------------------------------- */

/* Replaces the reference's whole timed host sequence cudaMalloc / H2D / gpu_matmul / D2H /
 * cudaFree (cuda_matmul.cu:135-165).  A, B, C host, row-major n*n.  kernel_s (may be NULL)
 * receives the device time of the multiply alone. */
int matmul_cuda(const float *A, const float *B, float *C, int n, int mode, double *kernel_s);

/* Replaces the launch gpu_matmul<<<grid, block>>>(d_A, d_B, d_C, nsize) (cuda_matmul.cu:155,
 * kernel :89-101): device pointers, C = A*B, asynchronous on `stream` (a cudaStream_t, NULL =
 * default stream).  Mode 3xTF32 keeps per-device scratch (split operand copies) that is
 * allocated on first use and grown on demand.  Calls on one device share that scratch: a call on a
 * different stream than the previous one first waits (on the GPU, through an event) for the previous
 * call's kernels, so concurrent streams and host threads get correct results but do not overlap. */
int matmul_cuda_device(const float *dA, const float *dB, float *dC, int n, int mode, void *stream);

/* Same multiply for callers that are not allowed to allocate device memory.  (No counterpart in the
 * reference, which cudaMallocs inside its timed window, cuda_matmul.cu:142-146.)  Think of a host
 * application that carved up the GPU's memory at start-up and forbids cudaMalloc afterwards, or a
 * latency-bound loop that cannot afford the device-wide synchronisation cudaMalloc/cudaFree imply.
 * This entry point calls no cudaMalloc, cudaFree or allocating library routine: its scratch is a fixed
 * 20 MiB array inside the library image, present on every GPU that loads libgauss_cuda.so until the
 * process exits.
 *   limit     n <= 1024; larger n returns GAUSS_CUDA_BAD_ARG (gauss_cuda_last_error() says why)
 *   set-up    the FIRST call on each device looks up the array's device address and sets kernel
 *             attributes; make that first call with the device current and before any stream capture
 *   sharing   all calls on a device use the same array, which keeps the previous call's contents;
 *             each call rewrites what it reads, so results do not depend on earlier calls, but two
 *             calls whose kernels can overlap on the GPU (different streams, no dependency) RACE on
 *             it -- serialise them
 * Arguments and result are those of matmul_cuda_device. */
int matmul_cuda_device_noalloc(const float *dA, const float *dB, float *dC, int n, int mode, void *stream);

/* ---- host-side input generators shared with the harness --------------------------------- */

/* initMatrix (gauss_internal_input.c:50-60): A[i][j] = 2*(min(i,j)+1), B[i] = i. */
void gauss_host_init_matrix(double *A, double *B, int n);
/* north_star's initialize_inputs(seed): srand(seed); A[i][j] = rand()/32768.0 row-major, the row's
 * absolute sum added to the diagonal; X_true[i] = i+1; B = A*X_true in double (the reference's
 * initRHS idea, gauss_openmp/gauss_external_input.c:88-106). */
void gauss_host_initialize_inputs(double *A, double *B, double *X_true, int n, unsigned seed);
/* cuda_matmul.cu:126-132: A[idx] = idx+1, B[idx] = 1/(idx+1). */
void gauss_host_matmul_fill(float *A, float *B, int n);
/* initMatrix(fname) of the external-input variant (gauss_external_input.c:32-84): "rows cols nnz" header, then
 * 1-based "row col value" lines up to a line with row 0.  Returns the order n (>0) and a malloc'd dense row-major
 * matrix in *A_out (release with gauss_host_free), or -1 when the file cannot be read (the reference exits). */
int  gauss_host_read_dat(const char *fname, double **A_out);
/* initRHS (gauss_external_input.c:88-106): X_pre[i] = i+1, R = A * X_pre. */
void gauss_host_init_rhs(const double *A, double *R, double *X_pre, int n);
void gauss_host_free(void *p);

/* ---- diagnostics ------------------------------------------------------------------------- */
const char *gauss_cuda_strerror(int status);
const char *gauss_cuda_last_error(void);     /* text of the last CUDA/NCCL failure in this thread */
int         gauss_cuda_device_count(void);   /* 0 when no usable CUDA device */

#ifdef __cplusplus
}
#endif
#endif /* GAUSS_CUDA_H */
