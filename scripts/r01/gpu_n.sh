#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pivot.py tests/test_gpu_gauss.py -q -m gpu -x --timeout 600 > gpurun_out/n_pytest.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/summary_n.txt
for st in 0 1; do
GAUSS_CUDA_TC_STATIC=$st timeout 300 python bench.py --workload gauss2048 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/n_bench_2048_static$st.json 2> gpurun_out/n_bench_2048_static$st.err; echo "bench2048 static=$st rc=$?" | tee -a gpurun_out/summary_n.txt
GAUSS_CUDA_TC_STATIC=$st timeout 300 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/n_bench_8192_static$st.json 2> gpurun_out/n_bench_8192_static$st.err; echo "bench8192 static=$st rc=$?" | tee -a gpurun_out/summary_n.txt
done
B2="python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$B2 > gpurun_out/n_plain2048.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/n_launches_2048.csv $B2 > gpurun_out/n_ncu2048.log 2>&1; echo "ncu list 2048 rc=$?" | tee -a gpurun_out/summary_n.txt
tail -n 4 gpurun_out/n_pytest.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/n_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and round(d['roofline'].get('share_of_step'),3), d['roofline']['avg_launch_ms'])
    except Exception as e: print(f, 'ERR', e)
PY
