#!/bin/bash
# CTA-pair kernel with the dynamic tile scheduler: parity + timing on one GPU
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_gauss.py -x -q -k "cta_pair or opt_in or at_scale" > gpurun_out/ah_pair_test.log 2>&1
rc=$?; echo "pair tests rc=$rc"; tail -5 gpurun_out/ah_pair_test.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ah_tests.log 2>&1
echo "full gpu suite rc=$?"; tail -3 gpurun_out/ah_tests.log
for st in 0 1; do
for w in gauss8192 gauss32768; do
  GAUSS_CUDA_TC_STATIC=$st timeout 600 python bench.py --workload $w --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ah_${w}_st$st.json 2> gpurun_out/ah_${w}_st$st.err
  echo "== TC_STATIC=$st $w"; grep '^{' gpurun_out/ah_${w}_st$st.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), round(d['roofline']['achieved'],1))"
done
done
