#!/bin/bash
mkdir -p gpurun_out
B8="python bench.py --workload gauss8192 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$B8 > gpurun_out/g_plain8192.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"diag_factor|urow|lcol" -s 30 -c 6 -o gpurun_out/prof_panel_v3 $B8 > gpurun_out/g_ncu_panel.log 2>&1; echo "ncu full panel rc=$?" | tee gpurun_out/summary_g.txt
ls -la gpurun_out/prof_panel_v3*
