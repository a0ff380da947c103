#!/bin/bash
# one GPU: ncu evidence (launch lists, DRAM traffic of the trailing update at the benchmarked outer block, full sets of the panel kernels)
mkdir -p gpurun_out/r2p
B="python bench.py --no-e2e --no-cpu-baseline --no-tf32-probe --no-secondary --steps 1 --warmup 1"
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
$B --workload gauss2048 > gpurun_out/r2p/plain2048.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2p/launches_gauss2048.csv $B --workload gauss2048 > gpurun_out/r2p/ncu2048.log 2>&1
echo "launch list 2048 rc=$?"
$B --workload gauss32768 > gpurun_out/r2p/plain32768.log 2>&1 &&
ncu --metrics $M --clock-control none -k regex:tc_gemm_pair -c 62 --csv --log-file gpurun_out/r2p/traffic_pair_gauss32768.csv $B --workload gauss32768 > gpurun_out/r2p/ncu_traffic.log 2>&1
echo "traffic rc=$?"
$B --workload gauss32768 > gpurun_out/r2p/plain32768b.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 600 --csv --log-file gpurun_out/r2p/launches_gauss32768.csv $B --workload gauss32768 > gpurun_out/r2p/ncu32768.log 2>&1
echo "launch list 32768 rc=$?"
$B --workload gauss8192 > gpurun_out/r2p/plain8192.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"diag_factor_pair|urow_lcol_mma|trsv_fused|tc_gemm_pair" -s 40 -c 12 -o gpurun_out/r2p/prof_panel_8192 -f $B --workload gauss8192 > gpurun_out/r2p/ncu_full.log 2>&1
echo "full set rc=$?"
GAUSS_CUDA_SCHED=cp $B --workload gauss8192 --nb 512 > gpurun_out/r2p/plain8192cp.log 2>&1 &&
GAUSS_CUDA_SCHED=cp ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 500 --csv --log-file gpurun_out/r2p/launches_gauss8192_cp.csv $B --workload gauss8192 --nb 512 > gpurun_out/r2p/ncu8192cp.log 2>&1
echo "launch list cp rc=$?"
ls -la gpurun_out/r2p | head -30
