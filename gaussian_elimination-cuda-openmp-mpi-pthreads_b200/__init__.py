"""Host-side mirror of the reference's Gauss / matmul interface over libgauss_cuda.so.

The reference program's interface is a handful of C functions on file-scope globals
(OpenMP_and_MPI/gauss_openmp/gauss_internal_input.c: allocate_memory :20-43, initMatrix
:50-60, computeGauss :118-138, solveGauss :144-159) and one kernel launch
(CUDA_and_OpenMP/Version-2/cuda_matmul.cu:155).  This module keeps those names and
argument meanings and forwards to the C ABI declared in include/gauss_cuda.h.

There is deliberately NO CPU path here: if the CUDA library is missing or no GPU is
present every compute call raises.  (The CPU oracle lives under oracle/ and is used
by tests/ and bench.py's cpu_baseline only.)
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libgauss_cuda.so")

MODE_3XTF32 = 0
MODE_FFMA = 1

OK, ZERO_PIVOT, BAD_ARG, CUDA_ERR, NCCL_ERR, NOT_READY = range(6)

# every symbol include/gauss_cuda.h declares (tests check the library exports them all)
ABI_SYMBOLS = (
    "gauss_cuda_init", "gauss_cuda_unique_id", "gauss_cuda_init_dist", "gauss_cuda_solve",
    "gauss_cuda_upload", "gauss_cuda_init_matrix_device", "gauss_cuda_solve_resident", "gauss_cuda_download",
    "gauss_cuda_compute_gauss", "gauss_cuda_solve_gauss", "gauss_cuda_download_factors",
    "gauss_cuda_set_pivoting", "gauss_cuda_get_pivots", "gauss_cuda_solve_refined", "gauss_cuda_free", "matmul_cuda", "matmul_cuda_device", "matmul_cuda_device_noalloc",
    "gauss_host_init_matrix", "gauss_host_initialize_inputs", "gauss_host_matmul_fill",
    "gauss_host_read_dat", "gauss_host_init_rhs", "gauss_host_free",
    "gauss_cuda_strerror", "gauss_cuda_last_error", "gauss_cuda_device_count",
)


class GaussStats(C.Structure):
    """gauss_cuda_stats (include/gauss_cuda.h)."""
    _fields_ = [
        ("h2d_s", C.c_double), ("factor_s", C.c_double), ("backsub_s", C.c_double), ("d2h_s", C.c_double),
        ("bcast_bytes", C.c_double), ("zero_pivot_row", C.c_int),
        ("n", C.c_int), ("n_padded", C.c_int), ("nb_outer", C.c_int), ("nb_inner", C.c_int),
        ("mode", C.c_int), ("rank", C.c_int), ("world", C.c_int),
        ("kernel_launches", C.c_longlong),
        ("update_s", C.c_double), ("update_flops", C.c_double), ("update_launches", C.c_int),
        ("overlapped_upload", C.c_int), ("graph_replayed", C.c_int),
    ]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RefineInfo(C.Structure):
    """gauss_cuda_refine_info (include/gauss_cuda.h)."""
    _fields_ = [
        ("iterations", C.c_int), ("converged", C.c_int), ("last_correction", C.c_double),
        ("residual_inf", C.c_double), ("factor_s", C.c_double), ("refine_s", C.c_double),
        ("kernel_launches", C.c_longlong),
    ]

    def asdict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class GaussCudaError(RuntimeError):
    def __init__(self, status, detail=""):
        self.status = status
        super().__init__(f"gauss_cuda status {status}: {detail}")


class SingularMatrix(GaussCudaError):
    """GAUSS_CUDA_ZERO_PIVOT - the reference prints 'The matrix is singular' and exits
    (gauss_internal_input.c:86-89)."""


_lib = None
_fp = C.POINTER(C.c_float)
_dp = C.POINTER(C.c_double)


def lib():
    """Loads libgauss_cuda.so; fails loudly when it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.gauss_cuda_init.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.c_int, C.c_int]
        L.gauss_cuda_unique_id.argtypes = [vp]
        L.gauss_cuda_init_dist.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int]
        L.gauss_cuda_solve.argtypes = [vp, _fp, _fp, _fp, _fp, C.POINTER(GaussStats)]
        L.gauss_cuda_upload.argtypes = [vp, _fp, _fp]
        L.gauss_cuda_init_matrix_device.argtypes = [vp]
        L.gauss_cuda_solve_resident.argtypes = [vp, C.POINTER(GaussStats)]
        L.gauss_cuda_compute_gauss.argtypes = [vp, C.POINTER(GaussStats)]
        L.gauss_cuda_solve_gauss.argtypes = [vp, C.POINTER(GaussStats)]
        L.gauss_cuda_download.argtypes = [vp, _fp, _fp]
        L.gauss_cuda_download_factors.argtypes = [vp, _fp]
        L.gauss_cuda_set_pivoting.argtypes = [vp, C.c_int]
        L.gauss_cuda_get_pivots.argtypes = [vp, C.POINTER(C.c_int)]
        L.gauss_cuda_solve_refined.argtypes = [vp, _dp, _dp, _dp, C.c_int, C.c_double, C.POINTER(RefineInfo)]
        L.gauss_host_read_dat.argtypes = [C.c_char_p, C.POINTER(_dp)]
        L.gauss_host_init_rhs.argtypes = [_dp, _dp, _dp, C.c_int]
        L.gauss_host_init_rhs.restype = None
        L.gauss_host_free.argtypes = [vp]
        L.gauss_host_free.restype = None
        L.gauss_cuda_free.argtypes = [vp]
        L.gauss_cuda_free.restype = None
        L.matmul_cuda.argtypes = [_fp, _fp, _fp, C.c_int, C.c_int, _dp]
        L.matmul_cuda_device.argtypes = [vp, vp, vp, C.c_int, C.c_int, vp]
        L.matmul_cuda_device_noalloc.argtypes = [vp, vp, vp, C.c_int, C.c_int, vp]
        L.gauss_host_init_matrix.argtypes = [_dp, _dp, C.c_int]
        L.gauss_host_init_matrix.restype = None
        L.gauss_host_initialize_inputs.argtypes = [_dp, _dp, _dp, C.c_int, C.c_uint]
        L.gauss_host_initialize_inputs.restype = None
        L.gauss_host_matmul_fill.argtypes = [_fp, _fp, C.c_int]
        L.gauss_host_matmul_fill.restype = None
        L.gauss_cuda_strerror.argtypes = [C.c_int]
        L.gauss_cuda_strerror.restype = C.c_char_p
        L.gauss_cuda_last_error.restype = C.c_char_p
        L.gauss_cuda_device_count.restype = C.c_int
        _lib = L
    return _lib


def _check(status):
    if status == OK:
        return
    L = lib()
    detail = L.gauss_cuda_strerror(status).decode()
    extra = L.gauss_cuda_last_error().decode()
    if status == ZERO_PIVOT:
        raise SingularMatrix(status, detail)
    raise GaussCudaError(status, f"{detail}; {extra}" if extra else detail)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def device_count():
    return lib().gauss_cuda_device_count()


# ---------------------------------------------------------------------------------------
# input generators (same names as the reference / north_star)
# ---------------------------------------------------------------------------------------
def initMatrix(nsize):
    """The reference's built-in fixture (gauss_internal_input.c:50-60): returns (matrix, B) in double."""
    A = np.empty((nsize, nsize), np.float64)
    B = np.empty(nsize, np.float64)
    lib().gauss_host_init_matrix(A.ctypes.data_as(_dp), B.ctypes.data_as(_dp), nsize)
    return A, B


def initialize_inputs(nsize, seed=1):
    """Seeded, diagonally dominant system with preset solution 1..n: returns (A, B, X_true) in double."""
    A = np.empty((nsize, nsize), np.float64)
    B = np.empty(nsize, np.float64)
    X = np.empty(nsize, np.float64)
    lib().gauss_host_initialize_inputs(A.ctypes.data_as(_dp), B.ctypes.data_as(_dp), X.ctypes.data_as(_dp), nsize, seed)
    return A, B, X


def read_dat(fname):
    """initMatrix(fname) of the external-input variant (gauss_external_input.c:32-84): dense double matrix."""
    p = _dp()
    n = lib().gauss_host_read_dat(os.fsencode(fname), C.byref(p))
    if n < 0:
        raise OSError("The matrix file open error")
    A = np.ctypeslib.as_array(p, shape=(n, n)).copy()
    lib().gauss_host_free(p)
    return A


def initRHS(A):
    """initRHS (gauss_external_input.c:88-106): returns (R, X__) with X__ = 1..n and R = A X__."""
    A = np.ascontiguousarray(A, np.float64)
    n = A.shape[0]
    R, Xp = np.empty(n), np.empty(n)
    lib().gauss_host_init_rhs(A.ctypes.data_as(_dp), R.ctypes.data_as(_dp), Xp.ctypes.data_as(_dp), n)
    return R, Xp


def matmul_fill(nsize):
    """cuda_matmul.cu:126-132: A[idx] = idx+1, B[idx] = 1/(idx+1)."""
    A = np.empty((nsize, nsize), np.float32)
    B = np.empty((nsize, nsize), np.float32)
    lib().gauss_host_matmul_fill(A.ctypes.data_as(_fp), B.ctypes.data_as(_fp), nsize)
    return A, B


# ---------------------------------------------------------------------------------------
# block-cyclic row ownership (host logic of the multi-GPU form; pure Python, testable on CPU)
# ---------------------------------------------------------------------------------------
def padded_order(n, inner=128):
    return (n + inner - 1) // inner * inner


def auto_outer_block(n, mode=MODE_3XTF32, world=1):
    """The outer block size the library picks for nb = 0 (ctx_create in csrc/gauss_cuda.cu)."""
    npad = padded_order(n)
    if mode == MODE_FFMA:
        nb = 256 if world > 1 else 128
    elif npad >= 24576 and world == 1:
        nb = 1024
    elif npad >= 8192:
        nb = 512
    elif npad >= 4096:
        nb = 256
    else:
        nb = 128
    return min(nb, npad)


def outer_blocks(n, nb):
    """[(row0, rows)] of the outer row blocks of the padded system."""
    npad = padded_order(n)
    nb = min(nb, npad)
    return [(r0, min(nb, npad - r0)) for r0 in range(0, npad, nb)]


def block_owner(block_index, world):
    """Row block r lives on rank r mod world (mirrors the per-worker row blocks of
    gauss_mpi/gauss_internal_input.c:140-155, made cyclic for load balance)."""
    return block_index % world


def local_row_map(n, nb, rank, world):
    """Global row index of every local row of `rank`, in local storage order."""
    rows = []
    for bi, (r0, cnt) in enumerate(outer_blocks(n, nb)):
        if block_owner(bi, world) == rank:
            rows.extend(range(r0, r0 + cnt))
    return np.asarray(rows, dtype=np.int64)


# ---------------------------------------------------------------------------------------
# the solver context
# ---------------------------------------------------------------------------------------
class GaussCuda:
    """One problem size on one GPU (or one rank of a multi-GPU group).

    allocate_memory(size) of the reference (gauss_internal_input.c:20-43) corresponds to
    constructing this object; gauss() = computeGauss + solveGauss.
    """

    def __init__(self, nsize, nb=0, mode=MODE_3XTF32, rank=0, world=1, device=None, nccl_id=None, ngpus=1):
        """ngpus > 1: ONE process drives GPUs 0..ngpus-1 (gauss_cuda_init's single-process group); rank/world/device/nccl_id
        describe one rank of the one-process-per-GPU form (gauss_cuda_init_dist)."""
        self._h = C.c_void_p()
        self.n = int(nsize)
        self.rank, self.world = rank, world
        L = lib()
        if ngpus > 1:
            self.world = ngpus
            _check(L.gauss_cuda_init(C.byref(self._h), self.n, ngpus, nb, mode))
        elif world == 1 and device is None:
            _check(L.gauss_cuda_init(C.byref(self._h), self.n, 1, nb, mode))
        else:
            idbuf = C.create_string_buffer(bytes(nccl_id), 128) if nccl_id is not None else None
            _check(L.gauss_cuda_init_dist(C.byref(self._h), self.n, rank, world, 0 if device is None else device,
                                          idbuf, nb, mode))
        self.stats = GaussStats()

    # -- host-buffer form --------------------------------------------------------------
    def gauss(self, A, B, want_b_after=False):
        """computeGauss + solveGauss on host data.  Returns X (and B after elimination)."""
        A32, B32 = _f32(A), _f32(B)
        if A32.shape != (self.n, self.n) or B32.shape != (self.n,):
            raise ValueError("shape mismatch")
        X = np.empty(self.n, np.float32)
        Bo = np.empty(self.n, np.float32) if want_b_after else None
        st = lib().gauss_cuda_solve(self._h, A32.ctypes.data_as(_fp), B32.ctypes.data_as(_fp), X.ctypes.data_as(_fp),
                                    Bo.ctypes.data_as(_fp) if want_b_after else None, C.byref(self.stats))
        _check(st)
        return (X, Bo) if want_b_after else X

    # -- device-resident form ----------------------------------------------------------
    def upload(self, A, B):
        A32, B32 = _f32(A), _f32(B)
        _check(lib().gauss_cuda_upload(self._h, A32.ctypes.data_as(_fp), B32.ctypes.data_as(_fp)))

    def initMatrix_device(self):
        """initMatrix (gauss_internal_input.c:50-60) generated on the GPU into the resident copy (no host data, no upload)."""
        _check(lib().gauss_cuda_init_matrix_device(self._h))

    def upload_ptr(self, a_ptr, b_ptr):
        """Same, from raw host addresses (e.g. pinned torch tensors): no copies made here."""
        _check(lib().gauss_cuda_upload(self._h, C.cast(a_ptr, _fp), C.cast(b_ptr, _fp)))

    def solve_ptr(self, a_ptr, b_ptr, x_ptr):
        _check(lib().gauss_cuda_solve(self._h, C.cast(a_ptr, _fp), C.cast(b_ptr, _fp), C.cast(x_ptr, _fp), None,
                                      C.byref(self.stats)))

    def solve_resident(self):
        _check(lib().gauss_cuda_solve_resident(self._h, C.byref(self.stats)))

    def computeGauss(self):
        """Forward elimination of the resident system (gauss_internal_input.c:118-138)."""
        _check(lib().gauss_cuda_compute_gauss(self._h, C.byref(self.stats)))

    def solveGauss(self):
        """Back substitution (gauss_internal_input.c:144-159)."""
        _check(lib().gauss_cuda_solve_gauss(self._h, C.byref(self.stats)))

    def download(self, want_b_after=False):
        X = np.empty(self.n, np.float32)
        Bo = np.empty(self.n, np.float32) if want_b_after else None
        _check(lib().gauss_cuda_download(self._h, X.ctypes.data_as(_fp), Bo.ctypes.data_as(_fp) if want_b_after else None))
        return (X, Bo) if want_b_after else X

    def gauss_refined(self, A, B, max_iters=20, tol=1e-14):
        """gauss() in the reference's own precision: A, B double (double **matrix, *B,
        gauss_internal_input.c:15-16), X double.  FP32 elimination on the GPU followed by iterative
        refinement with double residuals (gauss_cuda_solve_refined).  Returns (X, RefineInfo)."""
        A = np.ascontiguousarray(A, dtype=np.float64)
        B = np.ascontiguousarray(B, dtype=np.float64)
        if A.shape != (self.n, self.n) or B.shape != (self.n,):
            raise ValueError("A must be n x n and B of length n")
        X = np.empty(self.n, np.float64)
        info = RefineInfo()
        _check(lib().gauss_cuda_solve_refined(self._h, A.ctypes.data_as(_dp), B.ctypes.data_as(_dp),
                                              X.ctypes.data_as(_dp), int(max_iters), float(tol), C.byref(info)))
        return X, info

    def set_pivoting(self, enabled=True):
        """Partial pivoting as in the reference's external-input variant (gauss_external_input.c:123-148)."""
        _check(lib().gauss_cuda_set_pivoting(self._h, 1 if enabled else 0))

    def pivots(self):
        piv = np.empty(self.n, np.int32)
        _check(lib().gauss_cuda_get_pivots(self._h, piv.ctypes.data_as(C.POINTER(C.c_int))))
        return piv

    def download_factors(self):
        LU = np.empty((self.n, self.n), np.float32)
        _check(lib().gauss_cuda_download_factors(self._h, LU.ctypes.data_as(_fp)))
        return LU

    def close(self):
        if self._h:
            lib().gauss_cuda_free(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def gauss(A, B, mode=MODE_3XTF32, nb=0, want_b_after=False, pivoting=False):
    """One-shot gauss(): allocate, eliminate, back-substitute, free."""
    with GaussCuda(len(B), nb=nb, mode=mode) as g:
        if pivoting:
            g.set_pivoting(True)
        return g.gauss(A, B, want_b_after=want_b_after)


# ---------------------------------------------------------------------------------------
# multi-GPU bootstrap (one process per GPU; torch.distributed is only the id courier)
# ---------------------------------------------------------------------------------------
def unique_id():
    buf = C.create_string_buffer(128)
    _check(lib().gauss_cuda_unique_id(buf))
    return buf.raw


def init_distributed(nsize, nb=0, mode=MODE_3XTF32, device=None):
    """Creates this rank's GaussCuda inside an initialised torch.distributed group: rank 0 draws the
    NCCL id, the group broadcasts its 128 bytes, every rank joins the communicator."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", rank))
    if world == 1:
        return GaussCuda(nsize, nb=nb, mode=mode, rank=0, world=1, device=device)
    payload = [unique_id() if rank == 0 else None]
    dist.broadcast_object_list(payload, src=0)
    return GaussCuda(nsize, nb=nb, mode=mode, rank=rank, world=world, device=device, nccl_id=payload[0])


# ---------------------------------------------------------------------------------------
# matmul
# ---------------------------------------------------------------------------------------
def matmul_cuda(A, B, mode=MODE_3XTF32):
    """Host arrays in, host array out: the reference's whole GPU sequence (cuda_matmul.cu:135-165).
    Returns (C, kernel_seconds)."""
    A32, B32 = _f32(A), _f32(B)
    n = A32.shape[0]
    Cm = np.empty((n, n), np.float32)
    ks = C.c_double(0.0)
    _check(lib().matmul_cuda(A32.ctypes.data_as(_fp), B32.ctypes.data_as(_fp), Cm.ctypes.data_as(_fp), n, mode, C.byref(ks)))
    return Cm, ks.value


def matmul_cuda_device(dA, dB, dC, n, mode=MODE_3XTF32, stream=None):
    """Device pointers (ints, e.g. torch.Tensor.data_ptr()): replaces gpu_matmul<<<...>>> (cuda_matmul.cu:155)."""
    _check(lib().matmul_cuda_device(C.c_void_p(dA), C.c_void_p(dB), C.c_void_p(dC), n, mode,
                                    C.c_void_p(stream) if stream else None))


NOALLOC_MAX_N = 1024


def matmul_cuda_device_noalloc(dA, dB, dC, n, mode=MODE_3XTF32, stream=None):
    """matmul_cuda_device for hosts that must not allocate device memory after start-up (fixed memory
    budgets, latency-bound loops): the scratch is a 20 MiB array built into the library, so n <= 1024.
    The first call per device does a one-time set-up; calls that may overlap on the GPU must be serialised
    by the caller because they share that array."""
    _check(lib().matmul_cuda_device_noalloc(C.c_void_p(dA), C.c_void_p(dB), C.c_void_p(dC), n, mode,
                                            C.c_void_p(stream) if stream else None))
