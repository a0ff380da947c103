#!/bin/bash
mkdir -p gpurun_out
NG=${NG:-4}
for mt in 1 2 3; do
GAUSS_CUDA_OVERLAP_TILES=$mt GAUSS_CUDA_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 2955$mt bench.py --gpus $NG --steps 2 --warmup 1 --no-e2e > gpurun_out/w_bench_mt$mt.json 2> gpurun_out/w_trace_mt$mt.err
echo "== max_tiles=$mt"; grep "gauss_cuda trace" gpurun_out/w_trace_mt$mt.err | tail -n 2; grep '^{' gpurun_out/w_bench_mt$mt.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],2))"
done
GAUSS_CUDA_DIST_FORK=0 GAUSS_CUDA_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29559 bench.py --gpus $NG --steps 2 --warmup 1 --no-e2e > gpurun_out/w_bench_nofork.json 2> gpurun_out/w_trace_nofork.err
echo "== no fork"; grep "gauss_cuda trace" gpurun_out/w_trace_nofork.err | tail -n 2; grep '^{' gpurun_out/w_bench_nofork.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],2))"
