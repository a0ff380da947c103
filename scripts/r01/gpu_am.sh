#!/bin/bash
# multi-GPU panel factorisation split + early diagonal-region broadcast (GAUSS_CUDA_WIDE_SPLIT=1 default) vs 0
mkdir -p gpurun_out
NG=${NG:-2}
if [ "$NG" = "2" ]; then
timeout 600 python -m pytest tests/test_gpu_dist.py -x -q > gpurun_out/am_tests_g$NG.log 2>&1
echo "dist tests on $NG GPUs rc=$?"; tail -3 gpurun_out/am_tests_g$NG.log
fi
for ws in ${WSS:-1 0}; do
GAUSS_CUDA_WIDE_SPLIT=$ws timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 2958$ws bench.py --gpus $NG --workload gauss32768 --steps 3 --warmup 3 --no-e2e > gpurun_out/am_32768_g${NG}_ws$ws.json 2> gpurun_out/am_32768_g${NG}_ws$ws.err
echo "== gauss32768 x$NG WIDE_SPLIT=$ws rc=$?"; grep '^{' gpurun_out/am_32768_g${NG}_ws$ws.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],2), 'err', d.get('max_rel_err'), d['clocks'])"
done
