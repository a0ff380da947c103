#!/bin/bash
# round 2, call C: critical-path schedule on ONE GPU (GAUSS_CUDA_SCHED=cp): parity vs oracle, then timing
mkdir -p gpurun_out/r2c
export GAUSS_CUDA_SCHED=cp
timeout 600 python - > gpurun_out/r2c/cp_parity.log 2>&1 <<'PY'
import sys, os, numpy as np
sys.path.insert(0, 'tests')
from _pkg import pkg
import oracle_lib as ol
for n, nb in ((128, 128), (300, 128), (700, 256), (1000, 128), (1500, 512), (2048, 512), (2100, 256), (4500, 512)):
    A, B, Xt = pkg.initialize_inputs(n, 3)
    try:
        with pkg.GaussCuda(n, nb=nb, mode=pkg.MODE_3XTF32) as g:
            X, Ba = g.gauss(A, B, want_b_after=True)
            LU = g.download_factors()
        Xref, Bref, Uref = (ol.gauss(A, B, nthreads=8) if n <= 2100 else (Xt, None, None))
        print(n, nb, "errX", ol.max_rel_error(X, Xref), "errB", ol.max_rel_error(Ba, Bref) if Bref is not None else None,
              "resid", ol.scaled_residual(A, B, X), "errU", (np.max(np.abs(np.triu(LU.astype(np.float64), 1) - np.triu(Uref, 1))) / np.max(np.abs(Uref))) if Uref is not None else None, flush=True)
    except Exception as e:
        print(n, nb, "FAILED", repr(e), flush=True)
for n in (256, 1000, 2048):
    A, B = pkg.initMatrix(n)
    X, Ba = pkg.gauss(A, B, mode=pkg.MODE_3XTF32, nb=256, want_b_after=True)
    w = np.zeros(n, np.float32); w[0], w[-1] = -0.5, 0.5
    print("fixture", n, "exact X", np.array_equal(X, w), "B ok", np.array_equal(Ba[1:], np.full(n - 1, 0.5, np.float32)) and Ba[0] == 0, flush=True)
PY
echo "parity rc=$?" | tee gpurun_out/r2c/summary.txt
tail -n 14 gpurun_out/r2c/cp_parity.log
for wl in gauss8192 gauss16384 gauss32768; do
  timeout 300 python bench.py --workload $wl --steps 5 --no-cpu-baseline --no-secondary --no-e2e --no-tf32-probe > gpurun_out/r2c/bench_cp_$wl.json 2> gpurun_out/r2c/bench_cp_$wl.err; echo "$wl cp rc=$?" | tee -a gpurun_out/r2c/summary.txt
done
timeout 300 python bench.py --workload gauss32768 --nb 1024 --steps 5 --no-cpu-baseline --no-secondary --no-e2e --no-tf32-probe > gpurun_out/r2c/bench_cp_gauss32768_nb1024.json 2> gpurun_out/r2c/bench_cp_nb1024.err; echo "32768 nb1024 cp rc=$?" | tee -a gpurun_out/r2c/summary.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2c/bench_cp_*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][0])
        print(f, round(d['value']), round(d['ms_per_step'], 3), 'ms nb', d['config']['outer_block'], d['parity']['ok'], d['parity']['scaled_residual'], d['breakdown_ms'])
    except Exception as e:
        print(f, 'ERR', e)
PY
