#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/i_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_i.txt
timeout 300 python bench.py --workload gauss2048 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/i_bench_2048.json 2> gpurun_out/i_bench_2048.err; echo "bench2048 rc=$?" | tee -a gpurun_out/summary_i.txt
timeout 600 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/i_bench_8192.json 2> gpurun_out/i_bench_8192.err; echo "bench8192 rc=$?" | tee -a gpurun_out/summary_i.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/i_bench_32768.json 2> gpurun_out/i_bench_32768.err; echo "bench32768 rc=$?" | tee -a gpurun_out/summary_i.txt
tail -n 12 gpurun_out/i_pytest.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/i_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d['breakdown_ms'], 'err', d['max_rel_err'], 'launches', d['gpu_launches'], 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and round(d['roofline']['share_of_step'],3), 'e2e', d['e2e'] and round(d['e2e']['value']))
    except Exception as e: print(f, 'ERR', e)
PY
tail -n 3 gpurun_out/i_bench_*.err
