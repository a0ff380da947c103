#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/o_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_o.txt
for g in 1 0; do
GAUSS_CUDA_GRAPH=$g timeout 300 python bench.py --workload gauss2048 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/o_bench_2048_graph$g.json 2> gpurun_out/o_bench_2048_graph$g.err; echo "bench2048 graph=$g rc=$?" | tee -a gpurun_out/summary_o.txt
done
timeout 600 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/o_bench_8192.json 2> gpurun_out/o_bench_8192.err; echo "bench8192 rc=$?" | tee -a gpurun_out/summary_o.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/o_bench_32768.json 2> gpurun_out/o_bench_32768.err; echo "bench32768 rc=$?" | tee -a gpurun_out/summary_o.txt
for n in 2 ; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 --workload gauss16384 > gpurun_out/o_bench_16384_g$n.json 2> gpurun_out/o_bench_16384_g$n.err; echo "bench 16384 x$n rc=$?" | tee -a gpurun_out/summary_o.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/o_bench_32768_g$n.json 2> gpurun_out/o_bench_32768_g$n.err; echo "bench 32768 x$n rc=$?" | tee -a gpurun_out/summary_o.txt
done
tail -n 6 gpurun_out/o_pytest.log | cut -c1-300
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/o_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and round(d['roofline'].get('share_of_step'),3), 'e2e', d.get('e2e') and round(d['e2e']['value']), 'launches', d['gpu_launches'])
    except Exception as e: print(f, 'ERR', e)
PY
tail -n 4 gpurun_out/o_bench_*g2.err | tail -12
