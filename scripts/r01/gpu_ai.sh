#!/bin/bash
# multi-GPU with the CTA-pair kernel (GAUSS_CUDA_DIST_PAIR=1 default) vs without
mkdir -p gpurun_out
NG=${NG:-2}
if [ "$NG" = "2" ]; then
timeout 600 python -m pytest tests/test_gpu_dist.py -x -q > gpurun_out/ai_tests_g$NG.log 2>&1
echo "dist tests on $NG GPUs rc=$?"; tail -3 gpurun_out/ai_tests_g$NG.log
fi
for dp in ${DPS:-1 0}; do
for w in ${WL:-gauss32768 gauss16384}; do
GAUSS_CUDA_DIST_PAIR=$dp timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 2957$dp bench.py --gpus $NG --workload $w --steps 3 --warmup 3 > gpurun_out/ai_${w}_g${NG}_dp$dp.json 2> gpurun_out/ai_${w}_g${NG}_dp$dp.err
echo "== $w x$NG DIST_PAIR=$dp rc=$?"; grep '^{' gpurun_out/ai_${w}_g${NG}_dp$dp.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'roof', round(d['roofline']['achieved'],1), d['clocks'])"
done
done
