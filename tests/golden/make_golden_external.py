"""Regenerates tests/golden/external_vectors.npz: inputs and outputs of the reference's EXTERNAL-input
program (OpenMP_and_MPI/gauss_openmp/gauss_external_input.c: coordinate files, partial pivoting).

Run in the build container:  python tests/golden/make_golden_external.py
The matrices come from the reference's matrices_dense/*.dat (Harwell-Boeing test matrices stored as
'row col value' triples); they are kept here as COO triples, the "Error:" value is what the unmodified
reference binary (oracle/_ref/gauss_external_input) prints for the file.
"""
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
DATS = "/root/reference/OpenMP_and_MPI/gauss_openmp/matrices_dense"

out = {}
for name in ("matrix_10", "jpwh_991", "orsreg_1", "sherman5", "saylr4", "sherman3", "memplus"):
    path = os.path.join(DATS, name + ".dat")
    with open(path) as f:
        n = int(f.readline().split()[0])
        rows, cols, vals = [], [], []
        for line in f:
            p = line.split()
            if not p or int(p[0]) == 0:
                break
            rows.append(int(p[0]) - 1); cols.append(int(p[1]) - 1); vals.append(float(p[2]))
    out[f"{name}_n"] = np.array(n)
    out[f"{name}_rows"] = np.array(rows, np.int32)
    out[f"{name}_cols"] = np.array(cols, np.int32)
    out[f"{name}_vals"] = np.array(vals, np.float64)
    res = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "gauss_external_input"), path, str(os.cpu_count() or 4)],
                         capture_output=True, text=True).stdout
    out[f"{name}_ref_seconds"] = np.array(float(re.search(r"Time:\s+(\S+) seconds", res).group(1)))
    print(name, n, res.strip().splitlines()[-2:], flush=True)
    out[f"{name}_ref_error"] = np.array(float(re.search(r"Error: (\S+)", res).group(1)))
    # the file itself, verbatim, for the small ones: the parser tests read it back through the harness/ABI
    if name in ("matrix_10",):
        out[f"{name}_text"] = np.frombuffer(open(path, "rb").read(), dtype=np.uint8)
np.savez_compressed(os.path.join(HERE, "external_vectors.npz"), **out)
print({k: (v.shape, v.dtype) for k, v in out.items() if k.endswith("_n") or k.endswith("error")})
