#!/bin/bash
# round 2, call D (NG GPUs): benches at NG ranks, critical-path schedule (with trace) against the round-1 look-ahead
NG=${NG:-2}
CFGS=${CFGS:-"gauss32768:cp:512 gauss32768:cp:1024 gauss32768:lookahead:512 gauss16384:cp:512 gauss16384:cp:256"}
mkdir -p gpurun_out/r2d
nvidia-smi nvlink -gt d -i 0 > gpurun_out/r2d/nvlink_raw_g$NG.txt 2>&1
for cfg in $CFGS; do
  IFS=: read wl sched nb <<< "$cfg"
  tag=${wl}_g${NG}_${sched}_nb$nb
  GAUSS_CUDA_TRACE=1 GAUSS_CUDA_SCHED=$sched timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $NG --workload $wl --nb $nb --steps 5 --warmup 3 --no-e2e --no-tf32-probe > gpurun_out/r2d/bench_$tag.json 2> gpurun_out/r2d/bench_$tag.err; echo "$tag rc=$?" | tee -a gpurun_out/r2d/summary_g$NG.txt
  grep "trace" gpurun_out/r2d/bench_$tag.err | tail -n $NG | cut -c1-330
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2d/bench_*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][0])
        print(f, round(d['value']), round(d['ms_per_step'], 3), 'ms nb', d['config']['outer_block'], d['parity']['ok'], d['parity']['scaled_residual'], d['breakdown_ms'], (d.get('nvlink') or {}).get('measured_rx_bytes_per_step_rank0'))
    except Exception as e:
        print(f, 'ERR', e)
PY
