#!/bin/bash
mkdir -p gpurun_out
NG=${NG:-4}
GAUSS_CUDA_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $NG --steps 2 --warmup 1 --no-e2e > gpurun_out/v_bench_32768_g$NG.json 2> gpurun_out/v_trace_32768_g$NG.err; echo "rc=$?"
grep "gauss_cuda trace" gpurun_out/v_trace_32768_g$NG.err | tail -n $((NG*1)) 
grep '^{' gpurun_out/v_bench_32768_g$NG.json | cut -c1-200
