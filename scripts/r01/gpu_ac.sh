#!/bin/bash
# CTA-pair (cta_group::2) trailing update: parity first (bounded), then timing A/B
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_gauss.py -x -q -k "cta_pair" > gpurun_out/ac_pair_test.log 2>&1
rc=$?; echo "pair test rc=$rc"; tail -15 gpurun_out/ac_pair_test.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ac_tests.log 2>&1
echo "full gpu suite rc=$?"; tail -3 gpurun_out/ac_tests.log
for p in 0 1; do
for w in gauss8192 gauss16384 gauss32768; do
  GAUSS_CUDA_TC_PAIR=$p timeout 600 python bench.py --workload $w --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ac_bench_${w}_p$p.json 2> gpurun_out/ac_bench_${w}_p$p.err
  echo "== TC_PAIR=$p $w"; grep '^{' gpurun_out/ac_bench_${w}_p$p.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), d['roofline'])"
done
done
