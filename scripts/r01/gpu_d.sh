#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/d_smi.txt
timeout 900 python -m pytest tests/test_gpu_dist.py -q -m gpu -x --timeout 300 > gpurun_out/d_dist.log 2>&1; echo "dist pytest rc=$?" | tee gpurun_out/summary_d.txt
for n in 2 ; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 --workload gauss16384 > gpurun_out/d_bench_16384_g$n.json 2> gpurun_out/d_bench_16384_g$n.err; echo "bench 16384 x$n rc=$?" | tee -a gpurun_out/summary_d.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/d_bench_32768_g$n.json 2> gpurun_out/d_bench_32768_g$n.err; echo "bench 32768 x$n rc=$?" | tee -a gpurun_out/summary_d.txt
done
timeout 600 python bench.py --steps 3 --warmup 3 --workload gauss16384 --no-cpu-baseline > gpurun_out/d_bench_16384_g1.json 2> gpurun_out/d_bench_16384_g1.err; echo "bench 16384 x1 rc=$?" | tee -a gpurun_out/summary_d.txt
tail -n 15 gpurun_out/d_dist.log; cat gpurun_out/d_bench_*.json; tail -n 6 gpurun_out/d_bench_*g2.err
