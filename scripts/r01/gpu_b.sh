#!/bin/bash
# Full GPU test suite, harness programs, first bench lines.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 -x > gpurun_out/b_pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/summary_b.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/b_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary_b.txt
( ./harness/gauss_main -s 2048 -v | head -4; ./harness/gauss_main -s 2048 -m 1 | tail -3; ./harness/gauss_main 2048 1; ./harness/gauss_main 1000 3; ./harness/matmul_main 2048 0 skip; ./harness/matmul_main 2048 1 skip; ./harness/matmul_main 1001 0 skip ) > gpurun_out/b_harness.log 2>&1; echo "harness rc=$?" | tee -a gpurun_out/summary_b.txt
timeout 300 python bench.py --workload gauss2048 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_bench_2048.json 2> gpurun_out/b_bench_2048.err; echo "bench2048 rc=$?" | tee -a gpurun_out/summary_b.txt
timeout 300 python bench.py --workload gauss2048 --steps 10 --warmup 3 --no-cpu-baseline --mode ffma > gpurun_out/b_bench_2048_ffma.json 2>> gpurun_out/b_bench_2048.err; echo "bench2048 ffma rc=$?" | tee -a gpurun_out/summary_b.txt
timeout 300 python bench.py --workload matmul2048 --steps 20 > gpurun_out/b_bench_mm.json 2> gpurun_out/b_bench_mm.err; echo "bench mm rc=$?" | tee -a gpurun_out/summary_b.txt
timeout 300 python bench.py --workload matmul2048 --steps 20 --mode ffma > gpurun_out/b_bench_mm_ffma.json 2>> gpurun_out/b_bench_mm.err; echo "bench mm ffma rc=$?" | tee -a gpurun_out/summary_b.txt
timeout 600 python bench.py --workload gauss8192 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/b_bench_8192.json 2> gpurun_out/b_bench_8192.err; echo "bench8192 rc=$?" | tee -a gpurun_out/summary_b.txt
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/b_bench_32768.json 2> gpurun_out/b_bench_32768.err; echo "bench32768 rc=$?" | tee -a gpurun_out/summary_b.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/b_bench_ref.json 2> gpurun_out/b_bench_ref.err; echo "bench ref rc=$?" | tee -a gpurun_out/summary_b.txt
tail -n 25 gpurun_out/b_pytest.log gpurun_out/b_smoke.log gpurun_out/b_harness.log
cat gpurun_out/b_bench_*.json
tail -n 5 gpurun_out/b_bench_*.err
