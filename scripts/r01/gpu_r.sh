#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_gauss.py tests/test_gpu_pivot.py -q -m gpu -x --timeout 600 > gpurun_out/r_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_r.txt
timeout 300 python bench.py --workload gauss2048 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r_bench_2048.json 2> gpurun_out/r_bench_2048.err; echo "bench2048 rc=$?" | tee -a gpurun_out/summary_r.txt
timeout 600 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r_bench_8192.json 2> gpurun_out/r_bench_8192.err; echo "bench8192 rc=$?" | tee -a gpurun_out/summary_r.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r_bench_32768.json 2> gpurun_out/r_bench_32768.err; echo "bench32768 rc=$?" | tee -a gpurun_out/summary_r.txt
timeout 300 python bench.py --workload matmul2048 --steps 20 > gpurun_out/r_bench_mm.json 2> gpurun_out/r_bench_mm.err; echo "bench mm rc=$?" | tee -a gpurun_out/summary_r.txt
B2="python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
export GAUSS_CUDA_GRAPH=0
$B2 > gpurun_out/r_plain2048.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r_launches_2048.csv $B2 > gpurun_out/r_ncu2048.log 2>&1; echo "ncu list 2048 rc=$?" | tee -a gpurun_out/summary_r.txt
tail -n 5 gpurun_out/r_pytest.log | cut -c1-300
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'roof', d['roofline'] and round(d['roofline']['frac'],3), 'cpu', d.get('cpu_baseline'))
    except Exception as e: print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
grep diag_factor gpurun_out/r_launches_2048.csv | head -3 | cut -c1-200
