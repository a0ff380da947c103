#!/bin/bash
# round-1 v10 evidence: DRAM traffic of the CTA-pair kernel, launch list of one N=32768 solve, default bench line
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 600 $B > gpurun_out/ad_plain.json 2> gpurun_out/ad_plain.err || { echo "plain run failed"; exit 0; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:tc_gemm_pair -s 60 -c 60 --csv --log-file gpurun_out/ad_pair_traffic_32768.csv $B > gpurun_out/ad_ncu1.log 2>&1
echo "ncu traffic rc=$?"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/ad_launches_32768.csv $B > gpurun_out/ad_ncu2.log 2>&1
echo "ncu launches rc=$?"
timeout 900 python bench.py > gpurun_out/ad_bench_default.json 2> gpurun_out/ad_bench_default.err
echo "default bench rc=$?"; cat gpurun_out/ad_bench_default.json
