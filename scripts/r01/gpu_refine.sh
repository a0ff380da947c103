#!/bin/bash
# f4 bring-up: parity tests of the double-precision entry point, then the harness at a few sizes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_refine.py tests/test_gpu_gauss.py tests/test_zz_no_alloc.py -x -q > gpurun_out/refine_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/refine_tests.log
tail -5 gpurun_out/refine_tests.log
for a in "-s 2048 -r" "2048 1 -r" "8192 1 -r" "8192 1" "16384 2 -r" "-s 16384 -r"; do
  echo "== gauss_main $a" >> gpurun_out/refine_harness.log
  timeout 600 ./harness/gauss_main $a 2>&1 | grep -v "^$" | tail -4 >> gpurun_out/refine_harness.log
done
cat gpurun_out/refine_harness.log
