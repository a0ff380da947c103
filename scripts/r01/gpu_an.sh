#!/bin/bash
# from how many 256x256 tiles should a launch use the CTA-pair kernel (default: two rounds = 148)
mkdir -p gpurun_out
for mt in 148 74 30 8; do
for w in gauss2048 gauss8192 gauss16384; do
  GAUSS_CUDA_TC_PAIR_MIN_TILES=$mt timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/an_${w}_mt$mt.json 2> gpurun_out/an_${w}_mt$mt.err
  echo "== MIN_TILES=$mt $w $(grep '^{' gpurun_out/an_${w}_mt$mt.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3))")"
done
done
