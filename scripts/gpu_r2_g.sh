#!/bin/bash
# NG GPUs: sweeps of one environment switch inside one launch (SWEEPS="workload:nb:KEY=v1,v2 ...")
NG=${NG:-8}
mkdir -p gpurun_out/r2g
i=0
for sw in $SWEEPS; do
  IFS=: read wl nb kv <<< "$sw"
  i=$((i+1))
  GAUSS_CUDA_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $NG --workload $wl --nb $nb --steps 4 --warmup 2 --no-tf32-probe --env-sweep "$kv" > gpurun_out/r2g/sweep_${i}_g$NG.json 2> gpurun_out/r2g/sweep_${i}_g$NG.err; echo "$sw rc=$?"
  grep "trace cp\] rank 0" gpurun_out/r2g/sweep_${i}_g$NG.err | cut -c1-330
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2g/sweep_${i}_g$NG.json') if l.startswith('{')][0])
for e in d['sweep']: print(d['config'], e['env'], round(e['value']), round(e['ms_per_step'],3), e['breakdown_ms']['factor'], e['breakdown_ms']['backsub'], e['parity']['ok'], e['update_tflops_per_launch'])
PY
done
