#!/bin/bash
mkdir -p gpurun_out
NG=${NG:-4}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $NG --steps 3 --warmup 3 > gpurun_out/u_bench_32768_g$NG.json 2> gpurun_out/u_bench_32768_g$NG.err; echo "bench 32768 x$NG rc=$?" | tee -a gpurun_out/summary_u.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $NG --steps 3 --warmup 3 --no-e2e --workload gauss16384 > gpurun_out/u_bench_16384_g$NG.json 2> gpurun_out/u_bench_16384_g$NG.err; echo "bench 16384 x$NG rc=$?" | tee -a gpurun_out/summary_u.txt
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/u_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d.get('breakdown_ms'), 'roof', d.get('roofline') and round(d['roofline']['frac'],3), 'e2e', d.get('e2e') and round(d['e2e']['value']))
    except Exception as e: print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-400:])
PY
wc -l gpurun_out/u_bench_ref_g$NG.json
