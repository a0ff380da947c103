#!/bin/bash
# round 2, call B (2 GPUs): multi-GPU tests with the empty-rank / ragged cases on the round-1 schedule
mkdir -p gpurun_out/r2b
timeout 900 python -m pytest tests/test_gpu_dist.py -x -q -m gpu > gpurun_out/r2b/pytest_dist.log 2>&1; echo "pytest dist rc=$?" | tee gpurun_out/r2b/summary.txt
tail -n 15 gpurun_out/r2b/pytest_dist.log
