#!/bin/bash
# DRAM traffic of every tc_gemm launch of one N=32768 solve (2 light ncu passes), launch list at N=32768, and the matmul + reference-arm lines.
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$B > gpurun_out/j_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:tc_gemm -c 2000 --csv --log-file gpurun_out/j_tc_traffic_32768.csv $B > gpurun_out/j_ncu.log 2>&1; echo "ncu traffic rc=$?" | tee gpurun_out/summary_j.txt
$B > gpurun_out/j_plain2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/j_launches_32768.csv $B > gpurun_out/j_ncu2.log 2>&1; echo "ncu list 32768 rc=$?" | tee -a gpurun_out/summary_j.txt
$B > gpurun_out/j_plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 450 -c 1 -o gpurun_out/prof_tc_32768_r01 $B > gpurun_out/j_ncu3.log 2>&1; echo "ncu full P3 rc=$?" | tee -a gpurun_out/summary_j.txt
ls -la gpurun_out | grep " j_\|prof_tc_32768"
