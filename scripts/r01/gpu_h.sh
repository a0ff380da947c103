#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/h_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_h.txt
timeout 300 python bench.py --workload gauss2048 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/h_bench_2048.json 2> gpurun_out/h_bench_2048.err; echo "bench2048 rc=$?" | tee -a gpurun_out/summary_h.txt
timeout 600 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/h_bench_8192.json 2> gpurun_out/h_bench_8192.err; echo "bench8192 rc=$?" | tee -a gpurun_out/summary_h.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/h_bench_32768.json 2> gpurun_out/h_bench_32768.err; echo "bench32768 rc=$?" | tee -a gpurun_out/summary_h.txt
for n in 2 ; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 --workload gauss16384 > gpurun_out/h_bench_16384_g$n.json 2> gpurun_out/h_bench_16384_g$n.err; echo "bench 16384 x$n rc=$?" | tee -a gpurun_out/summary_h.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/h_bench_32768_g$n.json 2> gpurun_out/h_bench_32768_g$n.err; echo "bench 32768 x$n rc=$?" | tee -a gpurun_out/summary_h.txt
done
B2="python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$B2 > gpurun_out/h_plain2048.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/h_launches_2048.csv $B2 > gpurun_out/h_ncu2048.log 2>&1; echo "ncu list 2048 rc=$?" | tee -a gpurun_out/summary_h.txt
tail -n 12 gpurun_out/h_pytest.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/h_bench_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d['breakdown_ms'], 'err', d['max_rel_err'], 'launches', d['gpu_launches'], 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and round(d['roofline']['share_of_step'],3), 'e2e', d['e2e'] and round(d['e2e']['value']))
    except Exception as e: print(f, 'ERR', e)
PY
tail -n 5 gpurun_out/h_bench_*g2.err
