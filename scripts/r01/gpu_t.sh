#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dist.py -q -m gpu -x --timeout 300 > gpurun_out/t_dist.log 2>&1; echo "dist pytest rc=$?" | tee gpurun_out/summary_t.txt
n=2
for f in 1 0; do
GAUSS_CUDA_DIST_FORK=$f timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$f bench.py --gpus $n --steps 3 --warmup 3 --no-e2e > gpurun_out/t_bench_32768_g${n}_fork$f.json 2> gpurun_out/t_bench_32768_g${n}_fork$f.err; echo "bench 32768 x$n fork=$f rc=$?" | tee -a gpurun_out/summary_t.txt
GAUSS_CUDA_DIST_FORK=$f timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$f bench.py --gpus $n --steps 3 --warmup 3 --no-e2e --workload gauss16384 > gpurun_out/t_bench_16384_g${n}_fork$f.json 2> gpurun_out/t_bench_16384_g${n}_fork$f.err; echo "bench 16384 x$n fork=$f rc=$?" | tee -a gpurun_out/summary_t.txt
done
tail -n 4 gpurun_out/t_dist.log | cut -c1-300
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/t_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'roof', d['roofline'] and round(d['roofline']['frac'],3), 'err', d['max_rel_err'])
    except Exception as e: print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-400:])
PY
