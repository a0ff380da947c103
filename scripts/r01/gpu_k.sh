#!/bin/bash
# 4-GPU scaling probe with the SM-reservation knob for the NCCL broadcast.
mkdir -p gpurun_out
NG=${NG:-4}
for rs in 0 8 16; do
GAUSS_CUDA_TC_RESERVE_SMS=$rs timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 2951$((rs/8)) bench.py --gpus $NG --steps 3 --warmup 3 --no-e2e > gpurun_out/k_bench_32768_g${NG}_rs$rs.json 2> gpurun_out/k_bench_32768_g${NG}_rs$rs.err; echo "bench 32768 x$NG rs=$rs rc=$?" | tee -a gpurun_out/summary_k.txt
done
GAUSS_CUDA_TC_RESERVE_SMS=8 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus $NG --steps 3 --warmup 3 --no-e2e --workload gauss16384 > gpurun_out/k_bench_16384_g${NG}_rs8.json 2> gpurun_out/k_bench_16384_g${NG}_rs8.err; echo "bench 16384 x$NG rc=$?" | tee -a gpurun_out/summary_k.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $NG --steps 3 --warmup 3 --no-e2e --nb 256 > gpurun_out/k_bench_32768_g${NG}_nb256.json 2> gpurun_out/k_bench_32768_g${NG}_nb256.err; echo "bench 32768 nb256 x$NG rc=$?" | tee -a gpurun_out/summary_k.txt
for v in 0 1; do
GAUSS_CUDA_TC_VARIANT=$v timeout 600 python bench.py --workload gauss16384 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/k_bench_16384_g1_variant$v.json 2> gpurun_out/k_bench_16384_g1_variant$v.err; echo "bench 16384 x1 variant $v rc=$?" | tee -a gpurun_out/summary_k.txt
done
timeout 900 python -m pytest tests/test_gpu_dist.py -q -m gpu -x --timeout 300 > gpurun_out/k_dist.log 2>&1; echo "dist pytest rc=$?" | tee -a gpurun_out/summary_k.txt
tail -n 4 gpurun_out/k_dist.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/k_bench_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][0]); print(f, round(d['value']), 'GF', round(d['ms_per_step'],3),'ms', d['breakdown_ms'], 'err', d['max_rel_err'], 'roof', d['roofline'] and round(d['roofline']['frac'],3), d['roofline'] and round(d['roofline']['share_of_step'],3))
    except Exception as e: print(f, 'ERR', e)
PY
tail -n 3 gpurun_out/k_bench_*.err | tail -20
