#!/bin/bash
# source-level ncu capture of the panel kernels (N=2048): where do the 24 / 13 / 10 us go
mkdir -p gpurun_out
timeout 300 python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ab_plain.json 2> gpurun_out/ab_plain.err || exit 1
GAUSS_CUDA_GRAPH=0 timeout 900 ncu --set full --import-source on --clock-control none -k regex:'diag_factor_pair|urow_mma|lcol_mma' -s 9 -c 3 -o gpurun_out/ab_panel python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ab_ncu.log 2>&1
ls -la gpurun_out/ab_panel.ncu-rep
