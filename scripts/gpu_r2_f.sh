#!/bin/bash
# one GPU: correctness of the cp schedule after a change (parity on a few sizes + dist-free trace)
mkdir -p gpurun_out/r2f
export GAUSS_CUDA_SCHED=cp
timeout 600 python - > gpurun_out/r2f/cp_parity.log 2>&1 <<'PY'
import sys, os, numpy as np
sys.path.insert(0, 'tests')
from _pkg import pkg
import oracle_lib as ol
for n, nb in ((128, 128), (300, 128), (700, 256), (1500, 512), (2048, 512), (2100, 256), (4500, 512)):
    A, B, Xt = pkg.initialize_inputs(n, 3)
    try:
        with pkg.GaussCuda(n, nb=nb, mode=pkg.MODE_3XTF32) as g:
            X, Ba = g.gauss(A, B, want_b_after=True)
        Xref, Bref, Uref = (ol.gauss(A, B, nthreads=8) if n <= 2100 else (Xt, None, None))
        print(n, nb, "errX", ol.max_rel_error(X, Xref), "errB", ol.max_rel_error(Ba, Bref) if Bref is not None else None, "resid", ol.scaled_residual(A, B, X), flush=True)
    except Exception as e:
        print(n, nb, "FAILED", repr(e), flush=True)
A, B = pkg.initMatrix(2048)
X = pkg.gauss(A, B, mode=pkg.MODE_3XTF32, nb=256)
w = np.zeros(2048, np.float32); w[0], w[-1] = -0.5, 0.5
print("fixture exact", np.array_equal(X, w), flush=True)
PY
echo "parity rc=$?"; cat gpurun_out/r2f/cp_parity.log | tail -9
CFGS=${CFGS:-"gauss8192:512 gauss16384:512 gauss32768:512 gauss32768:1024"}
for cfg in $CFGS; do
  IFS=: read wl nb <<< "$cfg"
  GAUSS_CUDA_TRACE=1 timeout 300 python bench.py --workload $wl --nb $nb --steps 3 --warmup 2 --no-cpu-baseline --no-secondary --no-e2e --no-tf32-probe > gpurun_out/r2f/bench_${wl}_nb$nb.json 2> gpurun_out/r2f/bench_${wl}_nb$nb.err; echo "$wl nb$nb rc=$?"
  grep "trace cp" gpurun_out/r2f/bench_${wl}_nb$nb.err | tail -1
  python -c "
import json;d=json.loads([l for l in open('gpurun_out/r2f/bench_${wl}_nb$nb.json') if l.startswith('{')][0]);print(round(d['value']),round(d['ms_per_step'],3),d['breakdown_ms'])"
done
