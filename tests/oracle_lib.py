"""ctypes access to the CHECKERS under oracle/ (test infrastructure only).

* ``oracle()``      -> oracle/liboracle.so, our CPU restatement (gauss_oracle.c)
* ``RefGauss``      -> oracle/_ref/libgauss_ref.so, the reference's own OpenMP Gauss
                       program compiled unmodified (globals + initMatrix/computeGauss/
                       solveGauss, gauss_internal_input.c:15-16,50-60,118-159)
* ``ref_matmul()``  -> oracle/_ref/libmatmul_ref.so, the reference's cuda_matmul.cu

Nothing here reads /root/reference at run time; the .so files are built by
``__graft_entry__.build()`` (oracle/Makefile) and travel with the repo.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)


def _ptr(a, ty):
    return a.ctypes.data_as(ty)


def build_oracle():
    subprocess.run(["make", "-C", ORACLE_DIR], check=True, stdout=subprocess.DEVNULL)


_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        lib = C.CDLL(path)
        lib.oracle_init_matrix.argtypes = [_dp, _dp, _ip, C.c_int]
        lib.oracle_initialize_inputs.argtypes = [_dp, _dp, _dp, C.c_int, C.c_uint]
        lib.oracle_get_pivot.argtypes = [_dp, _dp, _ip, C.c_int, C.c_int]
        lib.oracle_compute_gauss.argtypes = [_dp, _dp, _ip, C.c_int, C.c_int]
        lib.oracle_compute_gauss_steps.argtypes = [_dp, _dp, _ip, C.c_int, C.c_int, C.c_int, C.c_int]
        lib.oracle_solve_gauss.argtypes = [_dp, _dp, _dp, _dp, C.c_int]
        lib.oracle_gauss.argtypes = [_dp, _dp, _dp, C.c_int, C.c_int]
        lib.oracle_max_rel_error.argtypes = [_dp, _dp, C.c_int]
        lib.oracle_max_rel_error.restype = C.c_double
        lib.oracle_normwise_error.argtypes = [_dp, _dp, C.c_int]
        lib.oracle_normwise_error.restype = C.c_double
        lib.oracle_scaled_residual_f32.argtypes = [_fp, _fp, _fp, C.c_int]
        lib.oracle_scaled_residual_f32.restype = C.c_double
        lib.oracle_matmul_fill.argtypes = [_fp, _fp, C.c_int]
        lib.oracle_matmul.argtypes = [_fp, _fp, _fp, C.c_int, C.c_int, C.c_int]
        lib.oracle_matmul_f64.argtypes = [_fp, _fp, _dp, C.c_int]
        lib.oracle_max_rel_diff_f32.argtypes = [_fp, _fp, C.c_size_t]
        lib.oracle_max_rel_diff_f32.restype = C.c_double
        lib.oracle_read_dat.argtypes = [C.c_char_p, C.POINTER(_dp)]
        lib.oracle_init_rhs.argtypes = [_dp, _dp, _dp, C.c_int]
        lib.oracle_compute_gauss_pivot.argtypes = [_dp, _dp, _ip, C.c_int, C.c_int]
        lib.oracle_gauss_pivot.argtypes = [_dp, _dp, _dp, C.c_int, C.c_int]
        _oracle = lib
    return _oracle


# ---- numpy-level wrappers around the restatement --------------------------

def init_matrix(n):
    A = np.empty((n, n), np.float64)
    B = np.empty(n, np.float64)
    oracle().oracle_init_matrix(_ptr(A, _dp), _ptr(B, _dp), None, n)
    return A, B


def initialize_inputs(n, seed):
    A = np.empty((n, n), np.float64)
    B = np.empty(n, np.float64)
    X = np.empty(n, np.float64)
    oracle().oracle_initialize_inputs(_ptr(A, _dp), _ptr(B, _dp), _ptr(X, _dp), n, seed)
    return A, B, X


def gauss(A, B, nthreads=1):
    """computeGauss + solveGauss on copies; returns (X, B_after, U) in double."""
    A = np.array(A, np.float64, order="C")
    B = np.array(B, np.float64)
    n = B.shape[0]
    rc = oracle().oracle_compute_gauss(_ptr(A, _dp), _ptr(B, _dp), None, n, nthreads)
    if rc != 0:
        raise ZeroDivisionError("The matrix is singular")
    V = np.empty(n)
    X = np.empty(n)
    oracle().oracle_solve_gauss(_ptr(A, _dp), _ptr(B, _dp), _ptr(V, _dp), _ptr(X, _dp), n)
    return X, B, A


def max_rel_error(X, Xref):
    X = np.ascontiguousarray(X, np.float64)
    Xref = np.ascontiguousarray(Xref, np.float64)
    return oracle().oracle_max_rel_error(_ptr(X, _dp), _ptr(Xref, _dp), X.shape[0])


def normwise_error(X, Xref):
    X = np.ascontiguousarray(X, np.float64)
    Xref = np.ascontiguousarray(Xref, np.float64)
    return oracle().oracle_normwise_error(_ptr(X, _dp), _ptr(Xref, _dp), X.shape[0])


def scaled_residual(A32, B32, X32):
    A32 = np.ascontiguousarray(A32, np.float32)
    B32 = np.ascontiguousarray(B32, np.float32)
    X32 = np.ascontiguousarray(X32, np.float32)
    return oracle().oracle_scaled_residual_f32(_ptr(A32, _fp), _ptr(B32, _fp), _ptr(X32, _fp), B32.shape[0])


def matmul_fill(n):
    A = np.empty((n, n), np.float32)
    B = np.empty((n, n), np.float32)
    oracle().oracle_matmul_fill(_ptr(A, _fp), _ptr(B, _fp), n)
    return A, B


def matmul(A, B, fused=0, nthreads=0):
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    n = A.shape[0]
    Cm = np.empty((n, n), np.float32)
    oracle().oracle_matmul(_ptr(A, _fp), _ptr(B, _fp), _ptr(Cm, _fp), n, fused, nthreads)
    return Cm


def matmul_f64(A, B):
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    n = A.shape[0]
    Cm = np.empty((n, n), np.float64)
    oracle().oracle_matmul_f64(_ptr(A, _fp), _ptr(B, _fp), _ptr(Cm, _dp), n)
    return Cm


# ---- external-input variant (coordinate files, partial pivoting) ------------

def read_dat(path):
    """oracle_read_dat: the reference's initMatrix(fname) (gauss_external_input.c:32-84) -> dense double matrix."""
    p = _dp()
    n = oracle().oracle_read_dat(os.fsencode(path), C.byref(p))
    if n < 0:
        raise OSError("The matrix file open error")
    A = np.ctypeslib.as_array(p, shape=(n, n)).copy()
    C.CDLL(None).free(p)
    return A


def init_rhs(A):
    """initRHS (gauss_external_input.c:88-106): X__ = 1..n, R = A X__."""
    A = np.ascontiguousarray(A, np.float64)
    n = A.shape[0]
    R, Xp = np.empty(n), np.empty(n)
    oracle().oracle_init_rhs(_ptr(A, _dp), _ptr(R, _dp), _ptr(Xp, _dp), n)
    return R, Xp


def gauss_pivot(A, R, nthreads=1):
    """Partial-pivoting computeGauss + solveGauss (gauss_external_input.c:123-205) on copies; returns X."""
    A = np.array(A, np.float64, order="C")
    R = np.array(R, np.float64)
    n = R.shape[0]
    X = np.empty(n)
    rc = oracle().oracle_gauss_pivot(_ptr(A, _dp), _ptr(R, _dp), _ptr(X, _dp), n, nthreads)
    if rc != 0:
        raise ZeroDivisionError("The matrix is singular")
    return X


def coo_to_dense(npz, name):
    n = int(npz[f"{name}_n"])
    A = np.zeros((n, n))
    A[npz[f"{name}_rows"], npz[f"{name}_cols"]] = npz[f"{name}_vals"]       # later entries win, like the parser
    return A


# ---- the reference's own object code --------------------------------------

def ref_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libgauss_ref.so"))


class RefGauss:
    """One problem instance inside the reference program's globals.

    The reference keeps `double **matrix, *B, *V, *C; int *swap` at file scope
    (gauss_internal_input.c:15-16) and never frees them, so each instance leaks
    its buffers by design; keep n small in tests.
    """

    def __init__(self, n):
        self.lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libgauss_ref.so"))
        self.n = n
        for f in ("allocate_memory", "initMatrix", "computeGauss", "solveGauss"):
            getattr(self.lib, f).argtypes = [C.c_int]
            getattr(self.lib, f).restype = None
        self.lib.allocate_memory(n)
        rows = C.POINTER(_dp).in_dll(self.lib, "matrix")
        self.A = np.ctypeslib.as_array(rows[0], shape=(n, n))      # matrix[0] is the flat block
        self.B = np.ctypeslib.as_array(_dp.in_dll(self.lib, "B"), shape=(n,))
        self.C = np.ctypeslib.as_array(_dp.in_dll(self.lib, "C"), shape=(n,))
        self.swap = np.ctypeslib.as_array(_ip.in_dll(self.lib, "swap"), shape=(n,))

    def init_fixture(self):
        self.lib.initMatrix(self.n)

    def load(self, A, B):
        self.A[...] = A
        self.B[...] = B
        self.swap[...] = np.arange(self.n, dtype=np.int32)

    def gauss(self, nthreads=1):
        omp = C.CDLL("libgomp.so.1")
        omp.omp_set_num_threads(int(nthreads))
        self.lib.computeGauss(self.n)
        self.lib.solveGauss(self.n)
        return self.C.copy(), self.B.copy(), self.A.copy()


_refmm = None


def ref_matmul():
    global _refmm
    if _refmm is None:
        lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libmatmul_ref.so"))
        for f in ("ref_seq_matmul", "ref_omp_matmul"):
            getattr(lib, f).argtypes = [_fp, _fp, _fp, C.c_int]
            getattr(lib, f).restype = None
        lib.ref_gpu_matmul.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        lib.ref_gpu_matmul.restype = C.c_float
        _refmm = lib
    return _refmm


def ref_cpu_matmul(A, B, omp=True):
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    n = A.shape[0]
    Cm = np.empty((n, n), np.float32)
    fn = ref_matmul().ref_omp_matmul if omp else ref_matmul().ref_seq_matmul
    fn(_ptr(A, _fp), _ptr(B, _fp), _ptr(Cm, _fp), n)
    return Cm
