#!/bin/bash
# graded chunk schedule of the overlapped upload: parity + e2e
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gauss.py -x -q > gpurun_out/z_tests.log 2>&1
echo "test_gpu_gauss rc=$?"; tail -2 gpurun_out/z_tests.log
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/z_e2e_${w}_$tag.json 2> gpurun_out/z_e2e_${w}_$tag.err
  echo "== $w $tag"; grep '^{' gpurun_out/z_e2e_${w}_$tag.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2))"
}
run graded gauss32768 X=1
run graded gauss16384 X=1
run off gauss16384 GAUSS_CUDA_STREAM=0
run graded gauss8192 GAUSS_CUDA_STREAM_MIN_N=8192
run off gauss8192 GAUSS_CUDA_STREAM=0
