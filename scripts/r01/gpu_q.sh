#!/bin/bash
# Final round-1 profiles on small workloads (ncu full captures are cheap here): one launch of every kernel class.
mkdir -p gpurun_out
B8="python bench.py --workload gauss8192 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
export GAUSS_CUDA_GRAPH=0
$B8 > gpurun_out/q_plain8192.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"diag_factor|urow|lcol|trsv_fused" -s 40 -c 4 -o gpurun_out/prof_panel_v5 $B8 > gpurun_out/q_ncu_panel.log 2>&1; echo "ncu panel rc=$?" | tee gpurun_out/summary_q.txt
$B8 > gpurun_out/q_plain8192b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:trsv_fused -c 1 -o gpurun_out/prof_trsv_v5 $B8 > gpurun_out/q_ncu_trsv.log 2>&1; echo "ncu trsv rc=$?" | tee -a gpurun_out/summary_q.txt
$B8 > gpurun_out/q_plain8192c.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 111 -c 1 -o gpurun_out/prof_tc_8192_v5 $B8 > gpurun_out/q_ncu_tc.log 2>&1; echo "ncu tc rc=$?" | tee -a gpurun_out/summary_q.txt
BM="python bench.py --workload matmul2048 --steps 3 --warmup 3"
$BM > gpurun_out/q_plainmm.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"tc_gemm|split_" -s 9 -c 3 -o gpurun_out/prof_matmul_v5 $BM > gpurun_out/q_ncu_mm.log 2>&1; echo "ncu matmul rc=$?" | tee -a gpurun_out/summary_q.txt
$B8 > gpurun_out/q_plain8192d.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/q_launches_8192.csv $B8 > gpurun_out/q_ncu8192.log 2>&1; echo "ncu list 8192 rc=$?" | tee -a gpurun_out/summary_q.txt
ls -la gpurun_out | grep "prof_.*v5\|q_launches"
