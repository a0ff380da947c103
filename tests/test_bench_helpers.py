"""bench.py's own checkers, exercised on the CPU: the parity block every bench line carries (north_star's two criteria,
evaluated per rank over its own rows and reduced with a max) must pass on a good solution, fail on a bad one, and give the
same verdict when the rows are split over several "ranks"; the tensor ceilings must be ordered as documented."""
import importlib.util
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)


def _system(n, seed=3):
    A = np.empty((n, n), np.float32)
    B = np.empty(n, np.float32)
    bench.fill_rows(A, B, n, 0, n, seed)
    return A, B, bench.preset_solution(n)


def test_parity_block_accepts_a_good_solution_and_rejects_a_bad_one():
    n = 700
    A, B, xt = _system(n)
    X = np.linalg.solve(A.astype(np.float64), B.astype(np.float64)).astype(np.float32)
    p = bench.parity_block(A, B, X, n, [(0, n)], xt, lambda v: v)
    assert p["ok"] and p["max_rel_err"] <= 1e-3 and p["scaled_residual"] <= 10.0
    assert p["tol"] == [1e-3, 10.0]
    Xbad = X.copy()
    Xbad[n // 2] *= 1.01                                   # 1 % off in one unknown: both criteria must notice
    pb = bench.parity_block(A, B, Xbad, n, [(0, n)], xt, lambda v: v)
    assert not pb["ok"] and pb["max_rel_err"] > 1e-3 and pb["scaled_residual"] > 10.0


def test_parity_block_is_the_same_when_rows_are_split_over_ranks():
    n = 640
    A, B, xt = _system(n, seed=5)
    X = np.linalg.solve(A.astype(np.float64), B.astype(np.float64)).astype(np.float32)
    whole = bench.parity_block(A, B, X, n, [(0, n)], xt, lambda v: v)
    # two "ranks" with interleaved row blocks, maxima reduced by hand exactly as dist.all_reduce(MAX) would
    parts = []
    for ranges in ([(0, 128), (256, 384), (512, 640)], [(128, 256), (384, 512)]):
        seen = []
        bench.parity_block(A, B, X, n, ranges, xt, lambda v: (seen.append(v), v)[1])
        parts.append(seen)                                 # [max |r_i|, max row sum] of that rank
    rmax, amax = max(p[0] for p in parts), max(p[1] for p in parts)
    resid = rmax / (amax * float(np.max(np.abs(X.astype(np.float64)))) * n * 2.0 ** -24)
    assert abs(resid - whole["scaled_residual"]) <= 1e-12 * max(1.0, whole["scaled_residual"])


def test_fill_rows_is_independent_of_the_row_partition():
    # every world size must see the same matrix: rows are seeded per 1024-row block
    n = 2304
    A1, B1, _ = _system(n, seed=1)
    A2 = np.empty_like(A1); B2 = np.empty_like(B1)
    for r0, r1 in ((0, 1024), (1024, 2048), (2048, 2304)):
        bench.fill_rows(A2, B2, n, r0, r1, 1)
    assert np.array_equal(A1, A2) and np.array_equal(B1, B2)
    assert np.all(np.abs(np.diag(A1)) > np.abs(A1).sum(axis=1) - np.abs(np.diag(A1)))   # strictly diagonally dominant


def test_tensor_ceilings_are_labelled_and_ordered():
    peaks = {"bf16_tflops": 1633.4, "bf16_tflops_sustained": 1392.7}
    c = bench.tensor_ceilings(peaks, {"tf32_tflops_burst": 744.0, "tf32_tflops_sustained": 619.0})
    assert abs(c["bf16_derived_sustained"] - 1392.7 / 6) < 1e-9 and abs(c["nominal"] - 375.0) < 1e-9
    assert c["tf32_measured_sustained"] < c["tf32_measured_burst"] < c["nominal"]
    assert "tf32_measured_burst" not in bench.tensor_ceilings(peaks, None)
