#!/bin/bash
# NG GPUs: the driver's own command line (default bench with e2e), then the N=16384 workload
NG=${NG:-2}
mkdir -p gpurun_out/r2q
for wl in gauss32768 gauss16384; do
  ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $NG --workload $wl --steps 5 --warmup 3 ) > gpurun_out/r2q/bench_${wl}_g$NG.json 2> gpurun_out/r2q/bench_${wl}_g$NG.err; echo "$wl g$NG rc=$?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2q/bench_*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][0])
        print(f, round(d['value']), round(d['ms_per_step'], 3), 'ms', d['parity']['ok'], 'e2e', d['e2e'] and (round(d['e2e']['ms_per_step'], 2), d['e2e']['host_memory'], d['e2e']['max_rel_err']), d['breakdown_ms'], d.get('nvlink', {}).get('algorithmic_rx_bytes_per_step_rank0'), d['roofline']['whole_step_frac'])
    except Exception as e:
        print(f, 'ERR', e)
PY
tail -3 gpurun_out/r2q/*.err | tail -12
