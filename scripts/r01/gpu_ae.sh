#!/bin/bash
# outer block size sweep with the v10 kernels
mkdir -p gpurun_out
run() { w=$1; nb=$2; timeout 600 python bench.py --workload $w --nb $nb --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ae_${w}_nb$nb.json 2> gpurun_out/ae_${w}_nb$nb.err
  echo "== $w nb=$nb"; grep '^{' gpurun_out/ae_${w}_nb$nb.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3))"; }
run gauss32768 256
run gauss32768 512
run gauss32768 768
run gauss32768 1024
run gauss16384 256
run gauss16384 512
run gauss16384 1024
run gauss8192 256
run gauss8192 512
run gauss2048 128
run gauss2048 256
run gauss2048 512
