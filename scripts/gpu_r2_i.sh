#!/bin/bash
# one GPU, cp schedule: parity, then a switch sweep (SWEEP="KEY=v1,v2") at two sizes
mkdir -p gpurun_out/r2i
export GAUSS_CUDA_SCHED=cp
CFGS="none" bash scripts/gpu_r2_f.sh 2>&1 | head -10
SWEEP=${SWEEP:-GAUSS_CUDA_CP_LOOK=1,0}
for wl in gauss8192 gauss16384; do
  GAUSS_CUDA_TRACE=1 timeout 300 python bench.py --workload $wl --nb 512 --steps 4 --warmup 2 --no-tf32-probe --env-sweep "$SWEEP" > gpurun_out/r2i/sweep_$wl.json 2> gpurun_out/r2i/sweep_$wl.err; echo "$wl rc=$?"
  grep "trace cp" gpurun_out/r2i/sweep_$wl.err | sed -n '3p;9p' | cut -c1-300
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2i/sweep_$wl.json') if l.startswith('{')][0])
for e in d['sweep']: print(d['config'], e['env'], round(e['value']), round(e['ms_per_step'],3), e['breakdown_ms']['factor'], e['parity'])
PY
done
