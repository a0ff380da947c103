#!/bin/bash
# First GPU bring-up: FFMA path, then the tcgen05 engine in isolated processes.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
export GAUSS_TEST_MODES=ffma
timeout 900 python -m pytest tests/test_gpu_gauss.py -q -m gpu --timeout 300 -x > gpurun_out/a_gauss_ffma.log 2>&1; echo "gauss ffma rc=$?" | tee -a gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_gpu_matmul.py -q -m gpu --timeout 300 -k "not variants" > gpurun_out/a_matmul_ffma.log 2>&1; echo "matmul ffma rc=$?" | tee -a gpurun_out/summary.txt
for v in 0 1; do
GAUSS_CUDA_TC_VARIANT=$v timeout 120 python - > gpurun_out/a_tc_v$v.log 2>&1 <<'PY'
import sys, numpy as np
sys.path.insert(0, 'tests')
from _pkg import pkg
import oracle_lib as ol
for n in (128, 256, 512, 2048):
    A, B = pkg.matmul_fill(n)
    C, ks = pkg.matmul_cuda(A, B, mode=pkg.MODE_3XTF32)
    C64 = ol.matmul_f64(A, B)
    err = np.max(np.abs(C - C64) / np.abs(C64))
    print("n", n, "max rel err", err, "kernel_s", ks, flush=True)
    if err > 1e-4:
        bad = np.argwhere(np.abs(C - C64) / np.abs(C64) > 1e-4)
        print("bad count", len(bad), "first", bad[:8].tolist(), "rows", sorted(set(bad[:,0].tolist()))[:20], "cols", sorted(set(bad[:,1].tolist()))[:40])
        print(C[:2,:8], C64[:2,:8])
rng = np.random.default_rng(0)
n = 512
A = rng.uniform(-1, 1, (n, n)).astype(np.float32); B = rng.uniform(-1, 1, (n, n)).astype(np.float32)
C, ks = pkg.matmul_cuda(A, B, mode=pkg.MODE_3XTF32)
C64 = A.astype(np.float64) @ B.astype(np.float64)
print("uniform err/bound", np.max(np.abs(C - C64) / (np.abs(A).astype(np.float64) @ np.abs(B).astype(np.float64))))
PY
echo "tc variant $v rc=$?" | tee -a gpurun_out/summary.txt
done
export GAUSS_TEST_MODES=tc
timeout 900 python -m pytest tests/test_gpu_gauss.py -q -m gpu --timeout 300 > gpurun_out/a_gauss_tc.log 2>&1; echo "gauss tc rc=$?" | tee -a gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_gpu_matmul.py -q -m gpu --timeout 300 > gpurun_out/a_matmul_tc.log 2>&1; echo "matmul tc rc=$?" | tee -a gpurun_out/summary.txt
tail -n 30 gpurun_out/a_*.log
