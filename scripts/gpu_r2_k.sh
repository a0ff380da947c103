#!/bin/bash
# one GPU: PDL along the single-GPU chain (bit-identity + timing at N = 2048 / 8192), with and without graph replay
mkdir -p gpurun_out/r2k
timeout 600 python - > gpurun_out/r2k/pdl_single.log 2>&1 <<'PY'
import os, sys, time, numpy as np
sys.path.insert(0, 'tests')
from _pkg import pkg
import oracle_lib as ol
def run(n, env):
    for k, v in env.items(): os.environ[k] = v
    A, B, Xt = pkg.initialize_inputs(n, 4)
    with pkg.GaussCuda(n, mode=pkg.MODE_3XTF32) as g:
        g.upload(A, B)
        for _ in range(5): g.solve_resident()
        t = []
        for _ in range(30):
            g.solve_resident(); st = g.stats; t.append((st.factor_s + st.backsub_s) * 1e3)
        X = g.download()
        gr = g.stats.graph_replayed
    for k in env: os.environ.pop(k)
    return X, float(np.median(t)), float(np.min(t)), gr
for n in (2048, 4096, 8192):
    X0, m0, b0, g0 = run(n, {})
    X1, m1, b1, g1 = run(n, {"GAUSS_CUDA_PDL_SINGLE": "1"})
    X2, m2, b2, g2 = run(n, {"GAUSS_CUDA_PDL_SINGLE": "1", "GAUSS_CUDA_GRAPH": "0"})
    X3, m3, b3, g3 = run(n, {"GAUSS_CUDA_GRAPH": "0"})
    print(n, "base graph", g0, round(m0, 4), round(b0, 4), "| pdl graph", g1, round(m1, 4), round(b1, 4), "| pdl plain", g2, round(m2, 4), round(b2, 4), "| plain", round(m3, 4),
          "| identical", np.array_equal(X0, X1), np.array_equal(X0, X2), flush=True)
PY
echo "rc=$?"; cat gpurun_out/r2k/pdl_single.log | tail -8
