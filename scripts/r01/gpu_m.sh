#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 -s > gpurun_out/m_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_m.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/m_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary_m.txt
timeout 300 python bench.py --workload gauss2048 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/m_bench_2048.json 2> gpurun_out/m_bench_2048.err; echo "bench2048 rc=$?" | tee -a gpurun_out/summary_m.txt
timeout 600 python bench.py --workload gauss16384 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/m_bench_16384.json 2> gpurun_out/m_bench_16384.err; echo "bench16384 rc=$?" | tee -a gpurun_out/summary_m.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/m_bench_32768.json 2> gpurun_out/m_bench_32768.err; echo "bench32768 rc=$?" | tee -a gpurun_out/summary_m.txt
timeout 300 python bench.py --workload matmul2048 --steps 20 > gpurun_out/m_bench_mm.json 2> gpurun_out/m_bench_mm.err; echo "bench mm rc=$?" | tee -a gpurun_out/summary_m.txt
( ./harness/gauss_main -s 2048 | tail -3; ./harness/gauss_main 2048 1 | tail -3 ) > gpurun_out/m_harness.log 2>&1
grep -E "passed|failed|Error \(GPU" gpurun_out/m_pytest.log | tail -12
tail -n 25 gpurun_out/m_pytest.log | cut -c1-300
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/m_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and d['roofline'].get('share_of_step'), 'e2e', d.get('e2e') and round(d['e2e']['value']), d.get('clocks'))
    except Exception as e: print(f, 'ERR', e)
PY
cat gpurun_out/m_harness.log
