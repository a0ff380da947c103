#!/bin/bash
# multi-GPU re-check with the v11 kernels: NG=2 runs the tests too
mkdir -p gpurun_out
NG=${NG:-2}
if [ "$NG" = "2" ]; then
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ag_tests_g$NG.log 2>&1
echo "gpu suite on $NG GPUs rc=$?"; tail -3 gpurun_out/ag_tests_g$NG.log
fi
for w in gauss32768 gauss16384; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $NG --workload $w --steps 3 --warmup 3 > gpurun_out/ag_${w}_g$NG.json 2> gpurun_out/ag_${w}_g$NG.err
echo "== $w x$NG rc=$?"; grep '^{' gpurun_out/ag_${w}_g$NG.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'roof', round(d['roofline']['achieved'],1), d['clocks'])"
done
