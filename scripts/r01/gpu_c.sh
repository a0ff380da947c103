#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/c_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee gpurun_out/summary_c.txt
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/c_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee -a gpurun_out/summary_c.txt
B8="python bench.py --workload gauss8192 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
B2="python bench.py --workload gauss2048 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$B8 > gpurun_out/c_plain8192.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/c_launches_8192.csv $B8 > gpurun_out/c_ncu8192.log 2>&1; echo "ncu list 8192 rc=$?" | tee -a gpurun_out/summary_c.txt
$B2 > gpurun_out/c_plain2048.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/c_launches_2048.csv $B2 > gpurun_out/c_ncu2048.log 2>&1; echo "ncu list 2048 rc=$?" | tee -a gpurun_out/summary_c.txt
$B8 > gpurun_out/c_plain8192b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 40 -c 8 -o gpurun_out/prof_tc_r01 $B8 > gpurun_out/c_ncu_tc.log 2>&1; echo "ncu full tc rc=$?" | tee -a gpurun_out/summary_c.txt
$B8 > gpurun_out/c_plain8192c.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"diag_factor|urow|lcol|trsv" -s 30 -c 8 -o gpurun_out/prof_panel_r01 $B8 > gpurun_out/c_ncu_panel.log 2>&1; echo "ncu full panel rc=$?" | tee -a gpurun_out/summary_c.txt
tail -n 8 gpurun_out/c_noalloc_alone.log gpurun_out/c_pytest.log
ls -la gpurun_out | tail -20
