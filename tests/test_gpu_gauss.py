"""GPU parity of gauss() (computeGauss + solveGauss) against the oracle, through the C ABI.

Tolerances are the north_star's: max |X_gpu - X_ref| / |X_ref| <= 1e-3 (norm-wise on the built-in
fixture, whose solution has n-2 exact zeros; element-wise in seeded mode) and scaled residual
||Ax-b||inf / (||A||inf ||x||inf n eps_fp32) <= 10.  The fixture is exact in FP32 and in TF32, so
there the comparison is bit-exact.
"""
import os

import numpy as np
import pytest

import oracle_lib as ol
from _pkg import pkg

pytestmark = pytest.mark.gpu
_sel = os.environ.get("GAUSS_TEST_MODES", "all")          # bring-up aid: ffma | tc | all
MODES = [m for m, tag in ((pkg.MODE_FFMA, "ffma"), (pkg.MODE_3XTF32, "tc")) if _sel in ("all", tag)]
TOL = 1e-3


def fixture_answer(n):
    x = np.zeros(n, np.float32)
    if n > 1:
        x[0], x[-1] = -0.5, 0.5
    b = np.full(n, 0.5, np.float32)
    b[0] = 0.0
    return x, b


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("n", [2, 8, 33, 128, 129, 300, 1000, 2048])
def test_fixture_known_answer_bit_exact(n, mode):
    # gauss_internal_input.c:45-48 - C = -0.5, 0, ..., 0, 0.5 and B = 0, 0.5, ... for ANY n
    A, B = pkg.initMatrix(n)
    X, Bafter = pkg.gauss(A, B, mode=mode, want_b_after=True)
    wx, wb = fixture_answer(n)
    assert np.array_equal(X, wx)
    assert np.array_equal(Bafter, wb)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("n", [5, 129, 1000, 2048])
def test_fixture_generated_on_the_device(n, mode):
    # gauss_cuda_init_matrix_device writes initMatrix's values (gauss_internal_input.c:50-60) with a kernel: the
    # result must be what the host fill + upload gives, bit for bit, and the known answer
    A, B = pkg.initMatrix(n)
    with pkg.GaussCuda(n, mode=mode) as g:
        g.initMatrix_device()
        g.solve_resident()
        X, Bafter = g.download(want_b_after=True)
        g.upload(A, B)
        g.solve_resident()
        X2, B2 = g.download(want_b_after=True)
    wx, wb = fixture_answer(n)
    assert np.array_equal(X, wx) and np.array_equal(Bafter, wb)
    assert np.array_equal(X, X2) and np.array_equal(Bafter, B2)


@pytest.mark.parametrize("mode", MODES)
def test_fixture_known_answer_at_scale(mode):
    # the same known answer holds for ANY n (gauss_internal_input.c:45-48) and the fixture is exact in FP32 and TF32:
    # a bit-exact check at a size where the CPU oracle would need minutes (3 outer blocks of 512, 64 inner blocks)
    n = 8192
    A, B = pkg.initMatrix(n)
    X, Bafter = pkg.gauss(A, B, mode=mode, want_b_after=True)
    wx, wb = fixture_answer(n)
    assert np.array_equal(X, wx)
    assert np.array_equal(Bafter, wb)


@pytest.mark.parametrize("mode", MODES)
def test_plain_launches_give_the_same_answer_as_graph_replay(mode, monkeypatch):
    # GAUSS_CUDA_GRAPH=0 (plain launches instead of graph replay) is read when the context is created
    n = 2100
    A, B, Xtrue = pkg.initialize_inputs(n, 12)
    with pkg.GaussCuda(n, mode=mode, nb=512) as g:
        g.upload(A, B)
        g.solve_resident()
        assert g.stats.graph_replayed == 1
        X0 = g.download()
    monkeypatch.setenv("GAUSS_CUDA_GRAPH", "0")
    with pkg.GaussCuda(n, mode=mode, nb=512) as g:
        g.upload(A, B)
        g.solve_resident()
        assert g.stats.graph_replayed == 0
        X2 = g.download()
    assert ol.max_rel_error(X0, Xtrue) <= TOL
    assert np.array_equal(X0, X2)                                # same kernels, same tile shapes, same order of sums


@pytest.mark.parametrize("mode", MODES)
def test_critical_path_schedule_on_one_gpu(mode, monkeypatch):
    # GAUSS_CUDA_SCHED=cp runs the multi-GPU schedule (diagonal region in the extended scratch, inverse-based panel GEMMs,
    # near / far updates) with a single rank: same tolerance against the oracle, and the fixture's exact answer
    monkeypatch.setenv("GAUSS_CUDA_SCHED", "cp")
    for n, nb in ((300, 128), (1500, 512), (2100, 256)):
        A, B, Xtrue = pkg.initialize_inputs(n, 3)
        Xref, Bref, _ = ol.gauss(A, B, nthreads=8)
        X, Bafter = pkg.gauss(A, B, mode=mode, nb=nb, want_b_after=True)
        assert ol.max_rel_error(X, Xref) <= TOL and ol.max_rel_error(Bafter, Bref) <= TOL
        assert ol.scaled_residual(A, B, X) <= 10.0
    A, B = pkg.initMatrix(2048)
    X, Bafter = pkg.gauss(A, B, mode=mode, nb=256, want_b_after=True)
    wx, wb = fixture_answer(2048)
    assert np.array_equal(X, wx) and np.array_equal(Bafter, wb)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("n,seed", [(64, 1), (200, 2), (512, 3), (1000, 1), (2048, 1)])
def test_seeded_matches_oracle(n, seed, mode):
    A, B, Xtrue = pkg.initialize_inputs(n, seed)
    Xref, Bref, Uref = ol.gauss(A, B, nthreads=8)
    with pkg.GaussCuda(n, mode=mode) as g:
        X, Bafter = g.gauss(A, B, want_b_after=True)
        assert g.stats.zero_pivot_row == -1 and g.stats.kernel_launches > 0
        LU = g.download_factors()
    assert ol.max_rel_error(X, Xref) <= TOL
    assert ol.max_rel_error(Bafter, Bref) <= TOL
    assert ol.scaled_residual(A, B, X) <= 10.0
    # unit-upper U: norm-wise against the oracle's (the reference zeroes L, gauss_internal_input.c:131)
    U = np.triu(LU.astype(np.float64), 1)
    assert np.max(np.abs(U - np.triu(Uref, 1))) <= TOL * np.max(np.abs(Uref))


@pytest.mark.parametrize("mode", MODES)
def test_golden_vectors(golden, mode):
    for n, seed in ((48, 1), (64, 2), (96, 3)):
        A, B = golden[f"seed{seed}_n{n}_A"], golden[f"seed{seed}_n{n}_B"]
        X = pkg.gauss(A, B, mode=mode)
        assert ol.max_rel_error(X, golden[f"seed{seed}_n{n}_X"]) <= TOL


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("nb", [128, 256, 512])
def test_outer_block_sizes_agree(nb, mode):
    n = 1100
    A, B, Xtrue = pkg.initialize_inputs(n, 5)
    X = pkg.gauss(A, B, mode=mode, nb=nb)
    assert ol.max_rel_error(X, Xtrue) <= TOL


@pytest.mark.parametrize("mode", MODES)
def test_zero_pivot_reported_as_singular(mode):
    # the reference would swap rows or print "The matrix is singular" (gauss_internal_input.c:75-97);
    # the GPU path reports the row instead of swapping (documented narrowing)
    n = 200
    A, B, _ = pkg.initialize_inputs(n, 3)
    A[:, 0] = 0.0
    with pkg.GaussCuda(n, mode=mode) as g:
        with pytest.raises(pkg.SingularMatrix):
            g.gauss(A, B)
        assert g.stats.zero_pivot_row == 0
    A, B, _ = pkg.initialize_inputs(n, 3)
    A[150, :] = A[149, :]                      # exact zero appears at pivot 150 during elimination
    with pkg.GaussCuda(n, mode=pkg.MODE_FFMA) as g:
        try:
            g.gauss(A, B)
            hit = g.stats.zero_pivot_row
        except pkg.SingularMatrix:
            hit = g.stats.zero_pivot_row
        assert hit in (-1, 150)                # rounding may leave a tiny non-zero pivot


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("n,nb,chunks", [(2100, 256, 3), (1500, 128, 8), (2048, 512, 2), (1100, 256, 64), (2100, 128, 0), (4000, 128, 0)])
def test_overlapped_upload_is_bit_identical(n, nb, chunks, mode, monkeypatch):
    # gauss_cuda_solve on a single GPU uploads the rows in chunks beside the elimination (by default from n = 16384;
    # forced here): every element must see the same updates in the same order as in the upload-first schedule
    A, B, Xtrue = pkg.initialize_inputs(n, 4)
    monkeypatch.setenv("GAUSS_CUDA_STREAM_MIN_N", "0")
    if chunks:                                 # 0: the default graded schedule (small chunks first)
        monkeypatch.setenv("GAUSS_CUDA_STREAM_CHUNKS", str(chunks))
    with pkg.GaussCuda(n, mode=mode, nb=nb) as g:
        X1, B1 = g.gauss(A, B, want_b_after=True)
        assert g.stats.overlapped_upload == 1
        LU1 = g.download_factors()
        X1b = g.gauss(A, B)                    # and again on the same context
        g.upload(A, B)
        g.solve_resident()
        assert g.stats.overlapped_upload == 0
        X2, B2 = g.download(want_b_after=True)
        LU2 = g.download_factors()
    monkeypatch.setenv("GAUSS_CUDA_STREAM", "0")
    with pkg.GaussCuda(n, mode=mode, nb=nb) as g:
        X3 = g.gauss(A, B)
        assert g.stats.overlapped_upload == 0
    assert ol.max_rel_error(X1, Xtrue) <= TOL
    assert np.array_equal(LU1, LU2)
    assert np.array_equal(X1, X2) and np.array_equal(X1, X1b) and np.array_equal(X1, X3) and np.array_equal(B1, B2)


@pytest.mark.parametrize("mode", MODES)
def test_auto_outer_block_matches_the_host_mirror(mode):
    # bench.py and multi-GPU callers place rows by pkg.auto_outer_block; it must be the library's own choice
    for n in (100, 2048, 4000, 4096, 8192):
        A, B = pkg.initMatrix(n)
        with pkg.GaussCuda(n, mode=mode) as g:
            g.gauss(A, B)
            assert g.stats.nb_outer == pkg.auto_outer_block(n, mode, 1)


def test_wide_split_of_the_row_panel_is_bit_identical(monkeypatch):
    # from n = 8192 the part of the row-panel factorisation right of the panel runs on its own stream, one inner block
    # behind the in-panel chain; same kernels and arithmetic, different launch ranges and order
    n = 8200                                   # np = 8320: 16 full outer blocks of 512 and a ragged one of 128
    A, B, Xtrue = pkg.initialize_inputs(n, 2)
    Xtrue = None
    with pkg.GaussCuda(n, mode=pkg.MODE_3XTF32) as g:
        X1 = g.gauss(A, B)
        g.upload(A, B)
        g.solve_resident()                     # graph replay of the same schedule
        X1r = g.download()
    monkeypatch.setenv("GAUSS_CUDA_WIDE_SPLIT", "0")
    with pkg.GaussCuda(n, mode=pkg.MODE_3XTF32) as g:
        X0 = g.gauss(A, B)
    assert ol.scaled_residual(A, B, X1) <= 10.0
    assert np.array_equal(X0, X1) and np.array_equal(X1, X1r)


def test_cta_pair_kernel_is_bit_identical(monkeypatch):
    # from ~150 tiles of 256 x 256 the trailing update runs on CTA pairs (tcgen05.mma.cta_group::2); every output
    # element must see the same MMAs in the same order as in the single-CTA kernel
    n = 4500                                   # np = 4608: ragged last tile row (18 row blocks of 256), 9 outer blocks
    A, B, Xtrue = pkg.initialize_inputs(n, 6)
    with pkg.GaussCuda(n, mode=pkg.MODE_3XTF32) as g:
        X1 = g.gauss(A, B)
        LU1 = g.download_factors()
    monkeypatch.setenv("GAUSS_CUDA_TC_PAIR", "0")
    with pkg.GaussCuda(n, mode=pkg.MODE_3XTF32) as g:
        X0 = g.gauss(A, B)
        LU0 = g.download_factors()
    assert ol.max_rel_error(X1, Xtrue) <= TOL
    assert np.array_equal(LU0, LU1) and np.array_equal(X0, X1)


@pytest.mark.parametrize("mode", MODES)
def test_overlapped_upload_reports_singular(mode, monkeypatch):
    monkeypatch.setenv("GAUSS_CUDA_STREAM_MIN_N", "0")
    n = 1000
    A, B, _ = pkg.initialize_inputs(n, 3)
    A[600:, :601] = 0.0                         # rows >= 600 are untouched by the first 600 steps and start with a zero
    with pkg.GaussCuda(n, mode=mode, nb=256) as g:
        with pytest.raises(pkg.SingularMatrix):
            g.gauss(A, B)
        assert g.stats.zero_pivot_row == 600 and g.stats.overlapped_upload == 1


@pytest.mark.parametrize("mode", MODES)
def test_resident_api_equals_host_api(mode):
    n = 700
    A, B, _ = pkg.initialize_inputs(n, 8)
    with pkg.GaussCuda(n, mode=mode) as g:
        X1 = g.gauss(A, B)
        g.upload(A, B)
        with pytest.raises(pkg.GaussCudaError):
            g.solveGauss()                     # nothing factored yet
        g.computeGauss()
        assert g.stats.factor_s > 0
        g.solveGauss()
        X2, B2 = g.download(want_b_after=True)
        g.solve_resident()                     # restores the pristine copy: repeatable
        X3 = g.download()
    assert np.array_equal(X1, X2) and np.array_equal(X2, X3)
    assert ol.max_rel_error(B2, ol.gauss(A, B)[1]) <= TOL


def test_solve_before_upload_is_rejected():
    with pkg.GaussCuda(64, mode=pkg.MODE_FFMA) as g:
        with pytest.raises(pkg.GaussCudaError) as e:
            g.solve_resident()
        assert e.value.status == pkg.NOT_READY


@pytest.mark.parametrize("mode", MODES)
def test_linearity_and_residual_at_scale(mode):
    # size-independent properties at a size the CPU oracle would take minutes for
    n = 8192
    rng = np.random.default_rng(11)
    A = rng.random((n, n), dtype=np.float32)
    A[np.arange(n), np.arange(n)] += A.sum(axis=1)
    # preset solution of bounded dynamic range: an ELEMENT-wise 1e-3 on x = 1..n would demand
    # |dx| <= 1e-3 next to entries of size n, i.e. 1e-7 * ||x||, which no FP32 solver can promise
    xt = 1.0 + (np.arange(n) % 8) * 0.125
    B = (A.astype(np.float64) @ xt).astype(np.float32)
    with pkg.GaussCuda(n, mode=mode) as g:
        X = g.gauss(A, B)
        X2 = g.gauss(A, 2.0 * B)
    print("n", n, "mode", mode, "max rel err", ol.max_rel_error(X, xt), "scaled residual", ol.scaled_residual(A, B, X))
    assert ol.max_rel_error(X, xt) <= TOL
    assert ol.scaled_residual(A, B, X) <= 10.0
    assert np.max(np.abs(X2 - 2.0 * X)) <= 1e-5 * np.max(np.abs(X2))       # scaling b by 2 is exact in binary FP


def test_north_star_generator_at_config4_size():
    # BASELINE config 4 (N = 16384) on the harness's own srand/rand initialize_inputs (diagonally dominant, x = 1..n):
    # the north_star's scaled residual ||Ax-b||inf / (||A||inf ||x||inf n eps) <= 10, and the solution norm-wise to
    # 1e-3 (element-wise 1e-3 next to entries of size n is below FP32 resolution at this size, see above);
    # the double-precision entry then has to reach the reference's accuracy element-wise
    n = 16384
    A, B, Xtrue = pkg.initialize_inputs(n, 1)
    A32, B32 = A.astype(np.float32), B.astype(np.float32)
    with pkg.GaussCuda(n, mode=pkg.MODE_3XTF32) as g:
        X = g.gauss(A32, B32)
        assert g.stats.zero_pivot_row == -1
    resid = ol.scaled_residual(A32, B32, X)
    nerr = ol.normwise_error(X, Xtrue)
    print("n", n, "scaled residual", resid, "norm-wise error", nerr)
    assert resid <= 10.0
    assert nerr <= TOL
