#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_gauss.py tests/test_gpu_pivot.py tests/test_gpu_matmul.py tests/test_zz_no_alloc.py -q -m gpu -x --timeout 600 > gpurun_out/p_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_p.txt
for ov in 1 0; do
GAUSS_CUDA_OVERLAP=$ov timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p_bench_32768_ov$ov.json 2> gpurun_out/p_bench_32768_ov$ov.err; echo "bench32768 ov=$ov rc=$?" | tee -a gpurun_out/summary_p.txt
GAUSS_CUDA_OVERLAP=$ov timeout 600 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p_bench_8192_ov$ov.json 2> gpurun_out/p_bench_8192_ov$ov.err; echo "bench8192 ov=$ov rc=$?" | tee -a gpurun_out/summary_p.txt
GAUSS_CUDA_OVERLAP=$ov timeout 300 python bench.py --workload gauss2048 --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p_bench_2048_ov$ov.json 2> gpurun_out/p_bench_2048_ov$ov.err; echo "bench2048 ov=$ov rc=$?" | tee -a gpurun_out/summary_p.txt
done
for mt in 3 12; do
GAUSS_CUDA_OVERLAP_TILES=$mt timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p_bench_32768_mt$mt.json 2> gpurun_out/p_bench_32768_mt$mt.err; echo "bench32768 mt=$mt rc=$?" | tee -a gpurun_out/summary_p.txt
done
GAUSS_CUDA_GRAPH=0 timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p_bench_32768_nograph.json 2> gpurun_out/p_bench_32768_nograph.err; echo "bench32768 nograph rc=$?" | tee -a gpurun_out/summary_p.txt
tail -n 6 gpurun_out/p_pytest.log | cut -c1-300
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/p_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'err', d['max_rel_err'], 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and round(d['roofline'].get('share_of_step'),3))
    except Exception as e: print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
PY
