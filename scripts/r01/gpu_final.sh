#!/bin/bash
# What the driver runs at round end: GPU tests, smoke, default bench, reference arm.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/final_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_final.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary_final.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err; echo "bench ref rc=$?" | tee -a gpurun_out/summary_final.txt
( time timeout 900 python bench.py ) > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?" | tee -a gpurun_out/summary_final.txt
timeout 300 python bench.py --workload gauss2048 --steps 20 > gpurun_out/final_bench_2048.json 2> gpurun_out/final_bench_2048.err; echo "bench2048 rc=$?" | tee -a gpurun_out/summary_final.txt
timeout 300 python bench.py --workload matmul2048 --steps 20 > gpurun_out/final_bench_mm.json 2> gpurun_out/final_bench_mm.err; echo "bench mm rc=$?" | tee -a gpurun_out/summary_final.txt
tail -n 4 gpurun_out/final_pytest.log; tail -n 2 gpurun_out/final_smoke.log; wc -l gpurun_out/final_bench.json; tail -n 5 gpurun_out/final_bench.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/final_bench*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), d['unit'], round(d['ms_per_step'],3),'ms', 'e2e', d.get('e2e') and round(d['e2e']['value']), 'roof', d.get('roofline') and (round(d['roofline']['frac'],3), d['roofline'].get('traffic')), 'cpu', d.get('cpu_baseline') and round(d['cpu_baseline']['value'],1), d.get('clocks'), 'launches', d.get('gpu_launches'))
    except Exception as e: print(f, 'ERR', e)
PY
