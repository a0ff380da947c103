#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_gauss.py tests/test_gpu_pivot.py -q -m gpu -x --timeout 600 > gpurun_out/s_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_s.txt
for f in 1 0; do
GAUSS_CUDA_INNER_FORK=$f timeout 300 python bench.py --workload gauss2048 --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/s_bench_2048_fork$f.json 2> gpurun_out/s_bench_2048_fork$f.err; echo "bench2048 fork=$f rc=$?" | tee -a gpurun_out/summary_s.txt
GAUSS_CUDA_INNER_FORK=$f timeout 600 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/s_bench_8192_fork$f.json 2> gpurun_out/s_bench_8192_fork$f.err; echo "bench8192 fork=$f rc=$?" | tee -a gpurun_out/summary_s.txt
GAUSS_CUDA_INNER_FORK=$f timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/s_bench_32768_fork$f.json 2> gpurun_out/s_bench_32768_fork$f.err; echo "bench32768 fork=$f rc=$?" | tee -a gpurun_out/summary_s.txt
done
tail -n 5 gpurun_out/s_pytest.log | cut -c1-300
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/s_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'roof', d['roofline'] and round(d['roofline']['frac'],3), 'err', d['max_rel_err'])
    except Exception as e: print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
