#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/f_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_f.txt
timeout 300 python bench.py --workload gauss2048 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/f_bench_2048.json 2> gpurun_out/f_bench_2048.err; echo "bench2048 rc=$?" | tee -a gpurun_out/summary_f.txt
timeout 300 python bench.py --workload gauss2048 --steps 10 --warmup 3 --no-cpu-baseline --mode ffma > gpurun_out/f_bench_2048_ffma.json 2>> gpurun_out/f_bench_2048.err; echo "bench2048 ffma rc=$?" | tee -a gpurun_out/summary_f.txt
timeout 600 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/f_bench_8192.json 2> gpurun_out/f_bench_8192.err; echo "bench8192 rc=$?" | tee -a gpurun_out/summary_f.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/f_bench_32768.json 2> gpurun_out/f_bench_32768.err; echo "bench32768 rc=$?" | tee -a gpurun_out/summary_f.txt
B2="python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$B2 > gpurun_out/f_plain2048.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/f_launches_2048.csv $B2 > gpurun_out/f_ncu2048.log 2>&1; echo "ncu list 2048 rc=$?" | tee -a gpurun_out/summary_f.txt
B8="python bench.py --workload gauss8192 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$B8 > gpurun_out/f_plain8192.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/f_launches_8192.csv $B8 > gpurun_out/f_ncu8192.log 2>&1; echo "ncu list 8192 rc=$?" | tee -a gpurun_out/summary_f.txt
tail -n 12 gpurun_out/f_pytest.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/f_bench_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d['breakdown_ms'], 'err', d['max_rel_err'], 'launches', d['gpu_launches'], 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and round(d['roofline']['share_of_step'],3))
    except Exception as e: print(f, 'ERR', e)
PY
