#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gauss.py tests/test_gpu_matmul.py -q -m gpu -x --timeout 600 > gpurun_out/l_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_l.txt
timeout 600 python bench.py --workload gauss16384 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/l_bench_16384.json 2> gpurun_out/l_bench_16384.err; echo "bench16384 rc=$?" | tee -a gpurun_out/summary_l.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/l_bench_32768.json 2> gpurun_out/l_bench_32768.err; echo "bench32768 rc=$?" | tee -a gpurun_out/summary_l.txt
timeout 300 python bench.py --workload matmul2048 --steps 20 > gpurun_out/l_bench_mm.json 2> gpurun_out/l_bench_mm.err; echo "bench mm rc=$?" | tee -a gpurun_out/summary_l.txt
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$B > gpurun_out/l_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:tc_gemm -s 444 -c 444 --csv --log-file gpurun_out/l_tc_traffic_32768.csv $B > gpurun_out/l_ncu.log 2>&1; echo "ncu traffic rc=$?" | tee -a gpurun_out/summary_l.txt
tail -n 3 gpurun_out/l_pytest.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/l_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and d['roofline'].get('share_of_step'), d.get('clocks'))
    except Exception as e: print(f, 'ERR', e)
PY
