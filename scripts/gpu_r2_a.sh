#!/bin/bash
# round 2, call A: full GPU suite with the new tests, smoke, default bench (parity + secondary + TF32 probe)
mkdir -p gpurun_out/r2a
nvidia-smi -L > gpurun_out/r2a/gpus.txt; nproc >> gpurun_out/r2a/gpus.txt; free -g >> gpurun_out/r2a/gpus.txt
timeout 1500 python -m pytest tests -x -q -m gpu -s > gpurun_out/r2a/pytest.log 2>&1; echo "pytest gpu rc=$?" | tee gpurun_out/r2a/summary.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/r2a/summary.txt
( time timeout 900 python bench.py ) > gpurun_out/r2a/bench.json 2> gpurun_out/r2a/bench.err; echo "bench rc=$?" | tee -a gpurun_out/r2a/summary.txt
tail -n 6 gpurun_out/r2a/pytest.log; tail -n 2 gpurun_out/r2a/smoke.log; tail -n 5 gpurun_out/r2a/bench.err
