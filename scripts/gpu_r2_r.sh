#!/bin/bash
# 4 GPUs: the driver's command line at N=4, and the single-process group through the C harness
mkdir -p gpurun_out/r2r
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 4 --steps 5 --warmup 3 ) > gpurun_out/r2r/bench_gauss32768_g4.json 2> gpurun_out/r2r/bench_gauss32768_g4.err; echo "bench g4 rc=$?"
python -c "
import json
d=json.loads([l for l in open('gpurun_out/r2r/bench_gauss32768_g4.json') if l.startswith('{')][0])
print(round(d['value']), round(d['ms_per_step'],3), d['parity']['ok'], d['parity']['scaled_residual'], 'e2e', round(d['e2e']['ms_per_step'],2), d['breakdown_ms'])"
./harness/gauss_main -s 4096 -g 4 -v > gpurun_out/r2r/harness_g4_fixture.log 2>&1; echo "harness -g 4 fixture rc=$?"; grep -E "Application|Known|GPU elim" gpurun_out/r2r/harness_g4_fixture.log
./harness/gauss_main -g 4 3000 5 > gpurun_out/r2r/harness_g4_seeded.log 2>&1; echo "harness -g 4 seeded rc=$?"; grep -E "Error|GPU elim" gpurun_out/r2r/harness_g4_seeded.log
./harness/gauss_main -g 3 -s 1000 -v | tail -n 2; echo "harness -g 3 rc=$?"
