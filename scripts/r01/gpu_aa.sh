#!/bin/bash
# tensor-core panel solves (GAUSS_CUDA_PANEL_MMA=1, default) vs the FFMA ones (=0): parity, timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/aa_tests.log 2>&1
echo "full gpu suite rc=$?"; tail -3 gpurun_out/aa_tests.log
for m in 0 1; do
for w in gauss2048 gauss8192 gauss32768; do
  GAUSS_CUDA_PANEL_MMA=$m timeout 600 python bench.py --workload $w --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/aa_bench_${w}_m$m.json 2> gpurun_out/aa_bench_${w}_m$m.err
  echo "== PANEL_MMA=$m $w"; grep '^{' gpurun_out/aa_bench_${w}_m$m.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3))"
done
done
for m in 0 1; do
GAUSS_CUDA_PANEL_MMA=$m timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/aa_launches_m$m.csv python bench.py --workload gauss8192 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/aa_ncu_m$m.log 2>&1
done
python - <<'PY'
import csv, collections
for d in (0, 1):
    rows = [r for r in csv.reader(open(f"gpurun_out/aa_launches_m{d}.csv", errors="ignore")) if len(r) > 10]
    hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value")
    acc = collections.defaultdict(list)
    for r in rows[1:]:
        try: acc[r[ki].split("(")[0]].append(float(r[vi].replace(",", "")))
        except ValueError: pass
    print("PANEL_MMA", d)
    for k, v in sorted(acc.items(), key=lambda kv: -sum(kv[1]))[:6]:
        print(f"  {k[:50]:50s} n={len(v):4d} total={sum(v)/1e3:9.1f} us avg={sum(v)/len(v)/1e3:7.2f} us max={max(v)/1e3:7.2f}")
PY
