"""Regenerates tests/golden/ref_vectors.npz from the REFERENCE'S OWN object code.

Run in the build container (needs oracle/_ref, i.e. /root/reference at build
time):   python tests/golden/make_golden.py

Every output below comes out of oracle/_ref/libgauss_ref.so (the unmodified
gauss_internal_input.c: initMatrix :50-60, computeGauss :118-138, solveGauss
:144-159) or oracle/_ref/libmatmul_ref.so (cuda_matmul.cu seq_matmul :28-39).
Seeded inputs are produced by the oracle's generator and stored alongside, so
the fixtures do not depend on the C library's rand() at test time.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle_lib as ol  # noqa: E402

out = {}
# built-in fixture, several sizes (known answer OI:45-48)
for n in (8, 16, 33):
    r = ol.RefGauss(n)
    r.init_fixture()
    X, Bafter, U = r.gauss(nthreads=2)
    out[f"fixture{n}_X"] = X
    out[f"fixture{n}_Bafter"] = Bafter
# seeded, diagonally dominant
for n, seed in ((48, 1), (64, 2), (96, 3)):
    A, B, Xtrue = ol.initialize_inputs(n, seed)
    r = ol.RefGauss(n)
    r.load(A, B)
    X, Bafter, U = r.gauss(nthreads=3)
    out[f"seed{seed}_n{n}_A"] = A
    out[f"seed{seed}_n{n}_B"] = B
    out[f"seed{seed}_n{n}_X"] = X
    out[f"seed{seed}_n{n}_Bafter"] = Bafter
    out[f"seed{seed}_n{n}_U"] = U
# zero leading pivot: exercises the reference's row-swap branch (OI:75-97)
n = 12
A, B, _ = ol.initialize_inputs(n, 7)
A[0, 0] = 0.0
A[5, 5] = 0.0
r = ol.RefGauss(n)
r.load(A, B)
X, Bafter, U = r.gauss(nthreads=1)
out["swap_n12_A"], out["swap_n12_B"], out["swap_n12_X"] = A, B, X
# matmul, reference CPU loop on the reference fill (CU2:126-132)
for n in (32, 65):
    A32, B32 = ol.matmul_fill(n)
    out[f"matmul{n}_C"] = ol.ref_cpu_matmul(A32, B32, omp=False)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_vectors.npz"), **out)
print("wrote", len(out), "arrays")
