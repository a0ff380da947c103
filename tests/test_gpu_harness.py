"""The drop-in programs themselves, run on the GPU: harness/gauss_main keeps the reference program's CLI, prints
and exit behaviour (OpenMP_and_MPI/gauss_openmp/gauss_internal_input.c:162-231) and reaches the GPU only through
the C ABI; harness/matmul_main does the same for cuda_matmul.cu:104-187.  The -v rows are compared with the output
of the reference's own program built with VERIFY on (oracle/_ref/gauss_verify), line for line."""
import os
import re
import subprocess

import numpy as np
import pytest

from _pkg import pkg

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GAUSS = os.path.join(ROOT, "harness", "gauss_main")
MATMUL = os.path.join(ROOT, "harness", "matmul_main")
REF_VERIFY = os.path.join(ROOT, "oracle", "_ref", "gauss_verify")
ROW = re.compile(r"^-?\d+\.\d{5} -?\d+\.\d{5}$")


def _run(cmd, timeout=600):
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)


def _rows(text):
    return [l for l in text.splitlines() if ROW.match(l.strip())]


@pytest.mark.skipif(not os.path.exists(GAUSS), reason="harness not built")
@pytest.mark.parametrize("extra", [[], ["-m", "1"], ["-d"]])
def test_reference_form_prints_the_reference_rows(extra):
    # gauss_main -s 2048 -v : banner, "Application time", then the VERIFY rows "%6.5f %5.5f" of B[i], C[i]
    # (gauss_internal_input.c:207, :223, :225-228); the rows must equal the reference program's own, which are
    # the analytic answer of :45-48
    n = 2048
    out = _run([GAUSS, "-s", str(n), "-t", "4", "-v"] + extra)
    assert out.returncode == 0, out.stderr
    assert f"Matrix Size: {n} ; Threads: " in out.stdout
    assert re.search(r"Application time: \d+\.\d+ Secs", out.stdout)
    rows = _rows(out.stdout)
    assert len(rows) == n
    if os.path.exists(REF_VERIFY):
        ref = _run([REF_VERIFY, "-s", str(n), "-t", "4"])
        assert ref.returncode == 0
        assert _rows(ref.stdout) == rows
    else:
        assert rows[0].split() == ["0.00000", "-0.50000"] and rows[-1].split() == ["0.50000", "0.50000"]
        assert all(r.split() == ["0.50000", "0.00000"] for r in rows[1:-1])


@pytest.mark.skipif(not os.path.exists(GAUSS), reason="harness not built")
@pytest.mark.parametrize("n,seed", [(2048, 1), (1000, 7)])
def test_seeded_form_meets_the_tolerance(n, seed):
    # north_star form `N [seed]`: srand/rand initialize_inputs, gauss(), the reference's error metric
    # (gauss_external_input.c:246-253) printed as "Error:"; exit status 0 only when it is <= 1e-3
    out = _run([GAUSS, str(n), str(seed)])
    assert out.returncode == 0, (out.stdout, out.stderr)
    m = re.search(r"Error: ([0-9.e+-]+)", out.stdout)
    assert m and float(m.group(1)) <= 1e-3
    assert f"Matrix Size: {n} ; Threads: " in out.stdout and "Application time:" in out.stdout


@pytest.mark.skipif(not os.path.exists(GAUSS), reason="harness not built")
def test_refined_form_reaches_double_accuracy():
    out = _run([GAUSS, "-r", "1500", "3"])
    assert out.returncode == 0, (out.stdout, out.stderr)
    m = re.search(r"Error: ([0-9.e+-]+)", out.stdout)
    assert m and float(m.group(1)) <= 1e-8


@pytest.mark.skipif(not os.path.exists(MATMUL), reason="harness not built")
def test_matmul_program_runs_and_prints_the_reference_lines():
    # cuda_matmul.cu prints "GPU Time: " (:166) (the CPU loops are skipped here: 2048^3 sequentially takes a minute)
    out = _run([MATMUL, "2048", "0", "skipcpu"])
    assert out.returncode == 0, (out.stdout, out.stderr)
    assert "GPU Time:" in out.stdout
    out = _run([MATMUL, "300", "1"])                     # small: the CPU twins run too and are compared
    assert out.returncode == 0, (out.stdout, out.stderr)
    assert "Seq (vectorized) Time:" in out.stdout and "OMP Time:" in out.stdout
