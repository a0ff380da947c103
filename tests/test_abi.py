"""CPU-side checks of the boundary: the library loads, exports every symbol the header declares,
refuses to compute without a GPU (no fallback), and its host-side helpers agree with the oracle."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
from _pkg import pkg, ROOT


def test_library_exports_every_declared_symbol():
    L = pkg.lib()
    header = open(os.path.join(ROOT, "include", "gauss_cuda.h")).read()
    declared = set(re.findall(r"\b((?:gauss_cuda|gauss_host|matmul_cuda)\w*)\s*\(", header))
    assert declared == set(pkg.ABI_SYMBOLS), declared ^ set(pkg.ABI_SYMBOLS)
    for name in declared:
        assert hasattr(L, name), name


def test_no_torch_types_or_oracle_in_product():
    # the product path must not route through the oracle or a CPU fallback
    pdir = os.path.join(ROOT, "gaussian_elimination-cuda-openmp-mpi-pthreads_b200")
    for dirpath, _, files in os.walk(pdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".c", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in text and "oracle_lib" not in text and "import oracle" not in text, f


def test_strerror_keeps_reference_message():
    L = pkg.lib()
    assert L.gauss_cuda_strerror(pkg.ZERO_PIVOT) == b"The matrix is singular"   # gauss_internal_input.c:87
    assert L.gauss_cuda_strerror(pkg.OK) == b"ok"


def test_bad_arguments_rejected():
    L = pkg.lib()
    h = C.c_void_p()
    assert L.gauss_cuda_init(C.byref(h), 0, 1, 0, 0) == pkg.BAD_ARG
    assert L.gauss_cuda_init(C.byref(h), 16, 1, 100, 0) == pkg.BAD_ARG      # nb not a multiple of 128
    assert L.gauss_cuda_init(C.byref(h), 16, 1, 0, 7) == pkg.BAD_ARG        # unknown mode
    assert L.gauss_cuda_init(C.byref(h), 16, 0, 0, 0) == pkg.BAD_ARG        # ngpus must be >= 1
    if pkg.device_count() < 2:                                              # a group of 2 needs 2 visible GPUs: loud failure, no fallback
        assert L.gauss_cuda_init(C.byref(h), 16, 2, 0, 0) == pkg.CUDA_ERR
        assert b"GPUs asked for" in L.gauss_cuda_last_error()
    assert L.gauss_cuda_solve(None, None, None, None, None, None) == pkg.BAD_ARG
    assert L.matmul_cuda(None, None, None, 4, 0, None) == pkg.BAD_ARG
    L.gauss_cuda_free(None)


@pytest.mark.skipif(pkg.device_count() > 0, reason="only meaningful without a GPU")
def test_fails_loudly_without_gpu():
    with pytest.raises(pkg.GaussCudaError) as e:
        pkg.gauss(np.eye(4), np.ones(4), mode=pkg.MODE_FFMA)
    assert e.value.status == pkg.CUDA_ERR
    a = np.ones((4, 4), np.float32)
    with pytest.raises(pkg.GaussCudaError):
        pkg.matmul_cuda(a, a, mode=pkg.MODE_FFMA)


def test_host_generators_match_oracle():
    A, B = pkg.initMatrix(37)
    Ao, Bo = ol.init_matrix(37)
    assert np.array_equal(A, Ao) and np.array_equal(B, Bo)
    A, B, X = pkg.initialize_inputs(50, seed=4)
    Ao, Bo, Xo = ol.initialize_inputs(50, 4)
    assert np.array_equal(A, Ao) and np.array_equal(B, Bo) and np.array_equal(X, Xo)
    Am, Bm = pkg.matmul_fill(33)
    Aom, Bom = ol.matmul_fill(33)
    assert np.array_equal(Am, Aom) and np.array_equal(Bm, Bom)


@pytest.mark.parametrize("n,nb,world", [(1000, 256, 4), (2048, 512, 8), (130, 128, 2), (128, 512, 3)])
def test_block_cyclic_ownership_partitions_rows(n, nb, world):
    npad = pkg.padded_order(n)
    seen = np.concatenate([pkg.local_row_map(n, nb, r, world) for r in range(world)])
    assert sorted(seen.tolist()) == list(range(npad))                       # every padded row exactly once
    blocks = pkg.outer_blocks(n, nb)
    assert sum(c for _, c in blocks) == npad and all(c % 128 == 0 for _, c in blocks)
    for r in range(world):
        rows = pkg.local_row_map(n, nb, r, world)
        assert np.all(np.diff(rows) > 0)                                    # local order = global order
        for g in rows[:: max(1, len(rows) // 7)]:
            assert pkg.block_owner(int(g) // min(nb, npad), world) == r


def test_auto_outer_block_rule():
    # host mirror of ctx_create's choice (the GPU suite checks it against stats.nb_outer); bench.py and multi-GPU callers
    # place rows with it, so it must be a multiple of the inner block, never exceed the padded order, and keep the
    # measured break points
    for n in (1, 100, 2048, 4000, 4096, 8191, 8192, 16384, 24576, 32768):
        for mode in (pkg.MODE_3XTF32, pkg.MODE_FFMA):
            for world in (1, 2, 8):
                nb = pkg.auto_outer_block(n, mode, world)
                assert nb % 128 == 0 and 128 <= nb <= pkg.padded_order(n)
    assert pkg.auto_outer_block(2048) == 128 and pkg.auto_outer_block(4096) == 256
    assert pkg.auto_outer_block(8192) == 512 and pkg.auto_outer_block(16384) == 512
    assert pkg.auto_outer_block(32768) == 1024 and pkg.auto_outer_block(32768, world=8) == 512
    assert pkg.auto_outer_block(32768, pkg.MODE_FFMA) == 128 and pkg.auto_outer_block(32768, pkg.MODE_FFMA, 4) == 256


def test_harness_cli_help_and_missing_gpu():
    exe = os.path.join(ROOT, "harness", "gauss_main")
    if not os.path.exists(exe):
        pytest.skip("harness not built")
    out = subprocess.run([exe, "-h"], capture_output=True, text=True)
    assert out.returncode == 0 and "Usage" in out.stdout
    if pkg.device_count() == 0:
        out = subprocess.run([exe, "-s", "64"], capture_output=True, text=True)
        assert out.returncode == 2 and "Matrix Size: 64" in out.stdout and "CUDA" in out.stderr
        # -g G (several GPUs from this one process): a group needs G visible devices -- loud failure, no fallback
        out = subprocess.run([exe, "-s", "64", "-g", "2"], capture_output=True, text=True)
        assert out.returncode == 2 and "GPUs asked for" in out.stderr
    out = subprocess.run([exe, "-s", "64", "-g", "0"], capture_output=True, text=True)
    assert "GPU count should be a natural number" in out.stderr


def test_external_harness_parses_its_arguments_without_gpu(tmp_path):
    # <matrixfile> [threads] [-r]: the reference's CLI (gauss_external_input.c:214-230) plus the refinement switch;
    # without a GPU the program must get as far as the device context and stop there with the library's message
    exe = os.path.join(ROOT, "harness", "gauss_external_main")
    if not os.path.exists(exe):
        pytest.skip("harness not built")
    z = np.load(os.path.join(ROOT, "tests", "golden", "external_vectors.npz"))
    path = tmp_path / "matrix_10.dat"
    path.write_bytes(z["matrix_10_text"].tobytes())
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode != 0 and "usage" in out.stderr
    out = subprocess.run([exe, str(path), "0"], capture_output=True, text=True)
    assert out.returncode != 0 and "natural number" in out.stderr
    if pkg.device_count() == 0:
        for args in ([str(path)], [str(path), "4"], [str(path), "4", "-r"], [str(path), "-r"]):
            out = subprocess.run([exe] + args, capture_output=True, text=True)
            assert out.returncode == 2 and "Matrix Size: 10" in out.stdout and "CUDA" in out.stderr, (args, out.stderr)

