#!/bin/bash
# blocked diagonal-block factor: parity, then A/B against the per-pivot-barrier kernel (GAUSS_CUDA_DIAG=0)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gauss.py tests/test_gpu_pivot.py tests/test_gpu_refine.py -x -q > gpurun_out/x_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/x_tests.log
tail -5 gpurun_out/x_tests.log
for d in 0 1; do
for w in gauss2048 gauss8192 gauss32768; do
  GAUSS_CUDA_DIAG=$d timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/x_bench_${w}_d$d.json 2> gpurun_out/x_bench_${w}_d$d.err
  echo "== DIAG=$d $w"; grep '^{' gpurun_out/x_bench_${w}_d$d.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3))"
done
done
for d in 0 1; do
GAUSS_CUDA_DIAG=$d timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/x_launches_d$d.csv python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/x_ncu_d$d.log 2>&1
done
python - <<'PY'
import csv, collections
for d in (0, 1):
    rows = [r for r in csv.reader(open(f"gpurun_out/x_launches_d{d}.csv", errors="ignore")) if len(r) > 10]
    hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value")
    acc = collections.defaultdict(list)
    for r in rows[1:]:
        try: acc[r[ki].split("(")[0]].append(float(r[vi].replace(",", "")))
        except ValueError: pass
    print("DIAG", d)
    for k, v in sorted(acc.items(), key=lambda kv: -sum(kv[1]))[:8]:
        print(f"  {k[:50]:50s} n={len(v):4d} total={sum(v)/1e3:9.1f} us avg={sum(v)/len(v)/1e3:7.2f} us")
PY
