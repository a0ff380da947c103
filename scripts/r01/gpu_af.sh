#!/bin/bash
# fused U-row + L-col launch, new automatic outer block: parity, timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/af_tests.log 2>&1
echo "full gpu suite rc=$?"; tail -3 gpurun_out/af_tests.log
for f in 0 1; do
for w in gauss2048 gauss8192; do
  GAUSS_CUDA_PANEL_FUSE=$f timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/af_${w}_f$f.json 2> gpurun_out/af_${w}_f$f.err
  echo "== FUSE=$f $w"; grep '^{' gpurun_out/af_${w}_f$f.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), d['config']['outer_block'])"
done
done
timeout 900 python bench.py --workload gauss16384 --no-cpu-baseline > gpurun_out/af_16384.json 2> gpurun_out/af_16384.err
grep '^{' gpurun_out/af_16384.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('16384', round(d['value'],1), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2))"
timeout 900 python bench.py > gpurun_out/af_bench_default.json 2> gpurun_out/af_bench_default.err
echo "default bench rc=$?"; grep '^{' gpurun_out/af_bench_default.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('32768', round(d['value'],1), round(d['ms_per_step'],3), 'nb', d['config']['outer_block'], 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), 'roof', round(d['roofline']['achieved'],1), d['clocks'])"
