#!/bin/bash
# one GPU: DRAM traffic of EVERY trailing-update launch of two N=32768 solves (the in-panel launches of the same kernel are captured too and told apart later)
mkdir -p gpurun_out/r2t
B="python bench.py --no-e2e --no-cpu-baseline --no-tf32-probe --no-secondary --steps 1 --warmup 1 --workload gauss32768"
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
$B > gpurun_out/r2t/plain.log 2>&1 &&
ncu --metrics $M --clock-control none -k regex:tc_gemm_pair -c 900 --csv --log-file gpurun_out/r2t/traffic_pair_gauss32768.csv $B > gpurun_out/r2t/ncu_traffic.log 2>&1
echo "traffic rc=$?"; wc -l gpurun_out/r2t/traffic_pair_gauss32768.csv
