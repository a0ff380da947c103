#!/bin/bash
# P1 split (in-panel chain on the main stream, wide part on its own stream): parity + timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gauss.py -x -q -k "wide_split or at_scale or overlapped" > gpurun_out/ak_split_test.log 2>&1
rc=$?; echo "split tests rc=$rc"; tail -6 gpurun_out/ak_split_test.log
if [ $rc -ne 0 ]; then exit 0; fi
for ws in 0 1; do
for w in gauss8192 gauss16384 gauss32768; do
  GAUSS_CUDA_WIDE_SPLIT=$ws timeout 600 python bench.py --workload $w --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ak_${w}_ws$ws.json 2> gpurun_out/ak_${w}_ws$ws.err
  echo "== WIDE_SPLIT=$ws $w"; grep '^{' gpurun_out/ak_${w}_ws$ws.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), d['clocks']['sm_mhz'])"
done
done
