#!/bin/bash
# diag-block factor with G pivots per barrier (GAUSS_CUDA_DIAG = 0 | 2 | 4 | 8): parity, bit-identity, timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/y_tests_all.log 2>&1
echo "full gpu suite rc=$?"; tail -3 gpurun_out/y_tests_all.log
for d in 4 8; do
GAUSS_CUDA_DIAG=$d timeout 900 python -m pytest tests/test_gpu_gauss.py tests/test_gpu_pivot.py -x -q > gpurun_out/y_tests_d$d.log 2>&1
echo "DIAG=$d pytest rc=$?"; tail -2 gpurun_out/y_tests_d$d.log
done
for d in 0 2 4 8; do
GAUSS_CUDA_DIAG=$d python - <<'PY'
import hashlib, os, sys
sys.path.insert(0, "tests")
from _pkg import pkg
import numpy as np
A, B, X = pkg.initialize_inputs(1500, 3)
for mode in (pkg.MODE_3XTF32, pkg.MODE_FFMA):
    with pkg.GaussCuda(1500, mode=mode) as g:
        x = g.gauss(A, B); LU = g.download_factors()
    print("DIAG", os.environ["GAUSS_CUDA_DIAG"], "mode", mode, hashlib.sha1(x.tobytes()).hexdigest()[:12], hashlib.sha1(LU.tobytes()).hexdigest()[:12])
PY
done
for d in 0 2 4 8; do
for w in gauss2048 gauss8192; do
  GAUSS_CUDA_DIAG=$d timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/y_bench_${w}_d$d.json 2> gpurun_out/y_bench_${w}_d$d.err
  echo "== DIAG=$d $w"; grep '^{' gpurun_out/y_bench_${w}_d$d.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3))"
done
done
for d in 2 4 8; do
GAUSS_CUDA_DIAG=$d timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/y_launches_d$d.csv python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/y_ncu_d$d.log 2>&1
done
python - <<'PY'
import csv, collections
for d in (2, 4, 8):
    rows = [r for r in csv.reader(open(f"gpurun_out/y_launches_d{d}.csv", errors="ignore")) if len(r) > 10]
    hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value")
    acc = collections.defaultdict(list)
    for r in rows[1:]:
        try: acc[r[ki].split("(")[0]].append(float(r[vi].replace(",", "")))
        except ValueError: pass
    print("DIAG", d)
    for k, v in sorted(acc.items(), key=lambda kv: -sum(kv[1]))[:4]:
        print(f"  {k[:50]:50s} n={len(v):4d} total={sum(v)/1e3:9.1f} us avg={sum(v)/len(v)/1e3:7.2f} us")
PY

# overlapped upload: e2e with and without, chunk counts
for cfg in "1 8" "0 8" "1 4" "1 16"; do
  set -- $cfg
  GAUSS_CUDA_STREAM=$1 GAUSS_CUDA_STREAM_CHUNKS=$2 timeout 900 python bench.py --workload gauss32768 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/y_e2e_32768_s$1_c$2.json 2> gpurun_out/y_e2e_32768_s$1_c$2.err
  echo "== 32768 STREAM=$1 CHUNKS=$2"; grep '^{' gpurun_out/y_e2e_32768_s$1_c$2.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', d['e2e'])"
done
for cfg in "1 8" "0 8"; do
  set -- $cfg
  for w in gauss16384 gauss8192; do
  GAUSS_CUDA_STREAM_MIN_N=8192 GAUSS_CUDA_STREAM=$1 GAUSS_CUDA_STREAM_CHUNKS=$2 timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/y_e2e_${w}_s$1.json 2> gpurun_out/y_e2e_${w}_s$1.err
  echo "== $w STREAM=$1"; grep '^{' gpurun_out/y_e2e_${w}_s$1.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', d['e2e'])"
  done
done
