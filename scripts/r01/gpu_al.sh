#!/bin/bash
# on-device fixture: parity + the harness's Application time with and without it
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gauss.py -x -q -k "generated_on_the_device or known_answer" > gpurun_out/al_test.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/al_test.log
for a in "-s 2048" "-s 2048 -d" "-s 16384" "-s 16384 -d" "-s 16384 -d" "-s 32768 -d"; do
  echo "== gauss_main $a" >> gpurun_out/al_harness.log
  timeout 600 ./harness/gauss_main $a 2>&1 | grep -v "^$" | tail -4 >> gpurun_out/al_harness.log
done
cat gpurun_out/al_harness.log
./harness/gauss_main -s 8 -d -v | head -12
