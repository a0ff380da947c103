#!/bin/bash
# NG GPUs: multi-GPU tests (world <= NG) then a P2P on/off sweep
NG=${NG:-2}
mkdir -p gpurun_out/r2h
timeout 1200 python -m pytest tests/test_gpu_dist.py -x -q -m gpu > gpurun_out/r2h/pytest_dist_g$NG.log 2>&1; echo "pytest dist rc=$?"
tail -n 6 gpurun_out/r2h/pytest_dist_g$NG.log
NG=$NG SWEEPS="${SWEEPS:-gauss32768:512:GAUSS_CUDA_P2P=1,0 gauss16384:512:GAUSS_CUDA_P2P=1,0}" bash scripts/gpu_r2_g.sh
