#!/usr/bin/env python
"""bench.py -- the headline benchmark: dense Gaussian elimination GFLOP/s (2/3 N^3) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload gauss32768|gauss16384|gauss2048|matmul2048]
                    [--mode tc|ffma] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one gauss() (computeGauss + solveGauss, gauss_internal_input.c:118-159) on a seeded,
diagonally dominant FP32 system.  `value` times K steps with the system resident in HBM (each step
restores the working copy device-to-device, eliminates, back-substitutes); `e2e` times the same call
through the host-buffer C-ABI entry (gauss_cuda_solve: H2D of [A|b] from pinned memory + solve + D2H
of X inside the timed region).  `roofline` describes the dominant kernel (the tcgen05 trailing update),
timed live with CUDA events on the library's stream.  `cpu_baseline` is the reference's own OpenMP
program (oracle/_ref, built from the unmodified source) timed on this box's host cores on a bounded
sample.  Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import re
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "gauss32768": ("gauss", 32768),
    "gauss16384": ("gauss", 16384),
    "gauss8192": ("gauss", 8192),
    "gauss2048": ("gauss", 2048),
    "matmul2048": ("matmul", 2048),
}
_REAL_STDOUT = None


def emit(line):
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


METRIC = "gauss_elim_gflops"
UNIT = "GFLOP/s"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def measure_tf32_peak(seconds=1.5):
    """Dense TF32 tensor-core rate of this GPU, measured outside the timed region with the library GEMM the roofline
    recipe uses for bf16 (torch.matmul, here on fp32 operands with TF32 allowed): 8192^3, best of 10 (burst) and back
    to back for `seconds` (sustained).  The 3xTF32 ceiling of the trailing update is this figure / 3."""
    import torch
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn((n, n), device="cuda", dtype=torch.float32)
        b = torch.randn((n, n), device="cuda", dtype=torch.float32)
        c = torch.empty_like(a)
        flop = 2.0 * n ** 3
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize()
        best = 0.0
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); torch.matmul(a, b, out=c); e1.record()
            torch.cuda.synchronize()
            best = max(best, flop / (e0.elapsed_time(e1) * 1e-3) * 1e-12)
        reps = max(10, int(seconds / (flop / (best * 1e12))))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        sustained = reps * flop / (e0.elapsed_time(e1) * 1e-3) * 1e-12
        del a, b, c
        torch.cuda.empty_cache()
        return {"tf32_tflops_burst": best, "tf32_tflops_sustained": sustained,
                "how": f"torch.matmul fp32 {n}^3 with allow_tf32 (cuBLAS TF32 tensor-core GEMM), best of 10 and {reps} back to back"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


# nominal dense TF32 rate of a B200 (half of the 2.25 PFLOP/s dense bf16 figure), for the fraction against the silicon limit
NOMINAL_TF32_TFLOPS = 1125.0


def tensor_ceilings(peaks, tf32):
    """3xTF32-effective ceilings (TFLOP/s of credited flops; every credited flop is three TF32 MMAs)."""
    out = {"bf16_derived_sustained": peaks["bf16_tflops_sustained"] / 6.0,
           "bf16_derived_burst": peaks.get("bf16_tflops", peaks["bf16_tflops_sustained"]) / 6.0,
           "nominal": NOMINAL_TF32_TFLOPS / 3.0}
    if tf32:
        out["tf32_measured_sustained"] = tf32["tf32_tflops_sustained"] / 3.0
        out["tf32_measured_burst"] = tf32["tf32_tflops_burst"] / 3.0
    return out


def parity_block(A, B, X, n, own_rows, xt, max_over_ranks):
    """north_star's two criteria on the solution the timed solves produced: max |X - x_true| / |x_true| <= 1e-3 and the
    scaled residual ||Ax-b||inf / (||A||inf ||x||inf n eps_fp32) <= 10, evaluated in double on the host from the FP32
    inputs; every rank does its own rows (A holds only those), the maxima are reduced over the ranks."""
    x64 = X.astype(np.float64)
    rmax = amax = 0.0
    for r0, r1 in own_rows:
        for s in range(r0, r1, 2048):
            e = min(r1, s + 2048)
            blk = A[s:e].astype(np.float64)
            rmax = max(rmax, float(np.max(np.abs(blk @ x64 - B[s:e].astype(np.float64)))))
            amax = max(amax, float(np.max(np.abs(blk).sum(axis=1))))
    rmax, amax = max_over_ranks(rmax), max_over_ranks(amax)
    eps = 2.0 ** -24
    resid = rmax / (amax * float(np.max(np.abs(x64))) * n * eps)
    err = float(np.max(np.abs(x64 - xt) / np.abs(xt)))
    return {"max_rel_err": err, "scaled_residual": resid, "tol": [1e-3, 10.0], "ok": bool(err <= 1e-3 and resid <= 10.0),
            "checked": "X of the last timed solve vs the preset solution; residual in double on the host over every rank's rows"}


# ------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        self.idx = gpu_index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.05)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------
# synthetic inputs: seeded per row block so every world size sees the same matrix
# ------------------------------------------------------------------------------------------
def fill_rows(A, B, n, r0, r1, seed):
    """Rows [r0, r1) of a diagonally dominant uniform matrix and of B = A * x_true (double accumulate)."""
    xt = preset_solution(n)
    step = 1024
    for s in range(r0, r1, step):
        e = min(r1, s + step)
        rng = np.random.default_rng([seed, s])
        blk = rng.random((e - s, n), dtype=np.float32)
        rows = np.arange(s, e)
        blk[np.arange(e - s), rows] += blk.sum(axis=1, dtype=np.float64).astype(np.float32)
        A[s:e] = blk
        B[s:e] = (blk.astype(np.float64) @ xt).astype(np.float32)


def preset_solution(n):
    """x_true of bounded dynamic range (1 .. 1.875): with the harness's x = 1..n an element-wise 1e-3 at
    n = 32768 would mean 3e-8 * ||x||, below FP32 resolution; the tolerance itself is unchanged."""
    return 1.0 + (np.arange(n) % 8) * 0.125


def _aligned_empty(shape, dtype=np.float32, align=4096):
    nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
    raw = np.empty(nbytes + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + nbytes].view(dtype).reshape(shape), raw


class HostBuffers:
    """A (n x n), B, X in host memory.  Only the row ranges this rank owns are ever touched, so only those are
    page-locked (cudaHostRegister per range): with 8 ranks each pins 1/8 of the matrix instead of all 4 GiB."""

    def __init__(self, n, row_ranges):
        import torch
        self.rt = torch.cuda.cudart()
        self.A, self._rawA = _aligned_empty((n, n))
        self.B, self._rawB = _aligned_empty((n,))
        self.X, self._rawX = _aligned_empty((n,))
        self.registered = []
        row_bytes = n * 4
        for r0, r1 in row_ranges:
            self._register(self.A.ctypes.data + r0 * row_bytes, (r1 - r0) * row_bytes)
        self._register(self.B.ctypes.data, n * 4)
        self._register(self.X.ctypes.data, n * 4)

    def _register(self, ptr, nbytes):
        lo = ptr & ~4095
        hi = (ptr + nbytes + 4095) & ~4095
        if any(lo < b and a < hi for a, b in self.registered):      # overlaps a page already pinned (small n)
            lo = max([lo] + [b for a, b in self.registered if a <= lo < b])
            if lo >= hi:
                return
        rc = self.rt.cudaHostRegister(lo, hi - lo, 0)
        if int(rc) == 0:
            self.registered.append((lo, hi))

    def close(self):
        for lo, _ in self.registered:
            self.rt.cudaHostUnregister(lo)
        self.registered = []


# ------------------------------------------------------------------------------------------
# CPU baseline: the reference's own program on this box's cores
# ------------------------------------------------------------------------------------------
def ref_binary():
    p = os.path.join(ROOT, "oracle", "_ref", "gauss_internal_input")
    return p if os.path.exists(p) else None


def run_reference_cpu(nsample, threads):
    """Runs the unmodified reference binary (gauss_internal_input.c main(), -s/-t CLI :171-205) and returns
    its own 'Application time' (window = initMatrix + computeGauss, :211-217)."""
    exe = ref_binary()
    if exe is None:
        return None
    out = subprocess.run([exe, "-s", str(nsample), "-t", str(threads)], capture_output=True, text=True, timeout=900)
    m = re.search(r"Application time: ([0-9.]+) Secs", out.stdout)
    return float(m.group(1)) if m else None


def cpu_baseline_gauss(cores, same_n=None):
    nsample = same_n if same_n else (6144 if cores >= 24 else 4096)
    t = run_reference_cpu(nsample, cores)
    if t is None:
        # the oracle port (oracle/gauss_oracle.c) stands in when the reference build did not travel
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        A, B = ol.init_matrix(nsample)
        t0 = time.perf_counter()
        ol.oracle().oracle_compute_gauss(A.ctypes.data_as(ol._dp), B.ctypes.data_as(ol._dp), None, nsample, cores)
        t = time.perf_counter() - t0
        kind = "port"
    else:
        kind = "reference"
    gf = 2.0 / 3.0 * nsample ** 3 / t * 1e-9
    out = {"value": gf, "unit": UNIT, "cores": cores, "kind": kind,
           "sample": f"full elimination of the built-in fixture at N={nsample} in double (gauss_internal_input -s {nsample} -t {cores}, {t:.3f} s"
                     + ("; the same N as the GPU arm)" if same_n else
                        f"; rate metric, the N=32768 run would take ~{2.0 / 3.0 * 32768 ** 3 / (gf * 1e9) / 60:.0f} min)")}
    if same_n and kind == "reference":       # configs[0]: the reference's sequential run (-t 1) beside its OpenMP run
        t1 = run_reference_cpu(nsample, 1)
        if t1:
            out["sequential"] = {"value": 2.0 / 3.0 * nsample ** 3 / t1 * 1e-9, "unit": UNIT, "cores": 1,
                                 "sample": f"gauss_internal_input -s {nsample} -t 1, {t1:.3f} s"}
    return out


def reference_arm(args, rank):
    """--impl reference: the reference's CPU implementation, all host threads, bounded sample per step."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    kind_w, n = WORKLOADS[args.workload]
    # the same N as our arm whenever the CPU finishes it in seconds (configs 1+2: N = 2048 on both sides); the
    # N = 16384 / 32768 workloads would take 1-10 minutes per step, so they are sampled as a rate at a smaller N
    nsample = n if n <= 8192 else (6144 if cores >= 24 else 4096)
    times = []
    for i in range(args.warmup + args.steps):
        t = run_reference_cpu(nsample, cores)
        if t is None:
            emit({"impl": "reference", "unavailable": "oracle/_ref/gauss_internal_input not built"})
            return
        if i >= args.warmup:
            times.append(t)
    t = sum(times) / len(times)
    gf = 2.0 / 3.0 * nsample ** 3 / t * 1e-9
    line = {"impl": "reference", "metric": METRIC, "value": gf, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic (reference initMatrix fixture)",
            "config": {"workload": args.workload, "n": n, "sample_n": nsample, "same_n": nsample == n},
            "cpu_baseline": {"value": gf, "unit": UNIT, "cores": cores, "kind": "reference",
                             "sample": f"gauss_internal_input -s {nsample} -t {cores}, its own Application time window"},
            "e2e": {"value": gf, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
class Dist:
    """Process-group plumbing (torch.distributed over NCCL when launched under torchrun)."""

    def __init__(self):
        import torch
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local_rank)
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
            self.dist = dist

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def _reduce(self, v, op):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, v):
        return self._reduce(v, self.dist.ReduceOp.MAX) if self.world > 1 else v

    def sum(self, v):
        return self._reduce(v, self.dist.ReduceOp.SUM) if self.world > 1 else v

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def nvlink_counters(gpu_index):
    """Cumulative NVLink data counters of one GPU in KiB (sum over links): (tx, rx) or None when the tool has none."""
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(gpu_index)], capture_output=True, text=True, timeout=20).stdout
    except (OSError, subprocess.TimeoutExpired):
        return None
    tx = sum(int(m) for m in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
    rx = sum(int(m) for m in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
    return (tx, rx) if (tx or rx) else None


def bench_gauss(args, pkg, D, workload, n, mode, peaks, peak_src, tf32, cores, steps, warmup, with_e2e=True, with_cpu=True,
                flush_l2=False, cpu_same_n=False):
    """One gauss workload: resident timing (value), parity, roofline of the trailing update, e2e, cpu baseline."""
    torch = D.torch
    rank, world, local_rank = D.rank, D.world, D.local_rank
    g = pkg.init_distributed(n, nb=args.nb, mode=mode, device=local_rank) if world > 1 else pkg.GaussCuda(n, nb=args.nb, mode=mode)
    probe_nb = args.nb if args.nb else pkg.auto_outer_block(n, mode, world)
    own_rows = [(r0, min(r0 + cnt, n)) for bi, (r0, cnt) in enumerate(pkg.outer_blocks(n, probe_nb))
                if pkg.block_owner(bi, world) == rank and r0 < n]
    hb = HostBuffers(n, own_rows)              # every rank page-locks its own rows only
    A, B, X = hb.A, hb.B, hb.X
    for r0, r1 in own_rows:
        fill_rows(A, B, n, r0, r1, args.seed)
    a_ptr, b_ptr, x_ptr = A.ctypes.data, B.ctypes.data, X.ctypes.data
    flop = 2.0 / 3.0 * n ** 3
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda") if flush_l2 else None

    # ---- resident (kernel) timing
    g.upload_ptr(a_ptr, b_ptr)
    for _ in range(max(warmup, 1)):             # at least one untimed solve (also fills the stats block)
        g.solve_resident()
    nb_outer = g.stats.nb_outer
    assert nb_outer == min(probe_nb, pkg.padded_order(n)), (nb_outer, probe_nb)
    sampler = ClockSampler(local_rank)
    nvl0 = nvlink_counters(local_rank) if world > 1 else None
    D.barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    dev_s = fac_s = bs_s = rest_s = upd_s = upd_flop = bcast_bytes = 0.0
    launches = upd_launches = graph_steps = 0
    for _ in range(steps):
        if flush is not None:
            flush.zero_()                        # outside the device-timed window (the library's own CUDA events)
        g.solve_resident()
        st = g.stats
        dev_s += st.h2d_s + st.factor_s + st.backsub_s
        rest_s += st.h2d_s; fac_s += st.factor_s; bs_s += st.backsub_s
        launches += st.kernel_launches
        upd_s += st.update_s; upd_flop += st.update_flops; upd_launches += st.update_launches
        bcast_bytes += st.bcast_bytes
        graph_steps += st.graph_replayed
    D.barrier()
    wall_s = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    nvl1 = nvlink_counters(local_rank) if world > 1 else None
    dev_s = D.max(dev_s)
    wall_s = D.max(wall_s)
    step_s = dev_s / steps
    value = flop / step_s * 1e-9
    Xres = g.download()
    xt = preset_solution(n)
    parity = parity_block(A, B, Xres, n, own_rows, xt, D.max)
    total_launches = int(D.sum(float(launches)))

    # ---- NVLink bytes of the panel broadcasts: counters of this GPU around the timed solves, per step
    nvlink = None
    if world > 1:
        rx_alg = bcast_bytes / steps
        nvlink = {"algorithmic_rx_bytes_per_step_rank0": rx_alg, "figure": "2 N^2 (G-1)/G bytes received per GPU and solve (SURVEY 8d: panel broadcast)",
                  "expected": 2.0 * n * n * (world - 1) / world}
        if nvl0 and nvl1:
            nvlink["measured_rx_bytes_per_step_rank0"] = (nvl1[1] - nvl0[1]) * 1024.0 / steps
            nvlink["measured_tx_bytes_per_step_rank0"] = (nvl1[0] - nvl0[0]) * 1024.0 / steps
            nvlink["how"] = "nvidia-smi nvlink -gt d on this rank's GPU before and after the timed solves (sum over links)"
        else:
            nvlink["measured_rx_bytes_per_step_rank0"] = None
            nvlink["how"] = "nvidia-smi exposes no NVLink data counters on this box; algorithmic figure only (library tally of received panel bytes)"

    # ---- end to end through the host-buffer entry point
    e2e = None
    if with_e2e:
        e2e_steps = max(1, min(steps, 3))
        g.solve_ptr(a_ptr, b_ptr, x_ptr)                     # warm
        D.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            g.solve_ptr(a_ptr, b_ptr, x_ptr)
        D.barrier()
        e2e_s = D.max(time.perf_counter() - t0) / e2e_steps
        h2d = D.sum(float(sum(r1 - r0 for r0, r1 in own_rows)) * n * 4 + n * 4)
        d2h = D.sum(float(n * 4))
        e2e = {"value": flop / e2e_s * 1e-9, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms_per_step": e2e_s * 1e3, "steps": e2e_steps, "host_memory": "pinned",
               "max_rel_err": float(np.max(np.abs(X.astype(np.float64) - xt) / xt))}

    # ---- roofline of the dominant kernel (rank 0's launches)
    ceil = tensor_ceilings(peaks, tf32)
    ffma_peak = 148 * 128 * 2 * ((clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)) * 1e6 * 1e-12
    roofline = None
    if upd_s > 0:
        ach = upd_flop / upd_s * 1e-12
        if mode == pkg.MODE_3XTF32:
            peak = ceil.get("tf32_measured_sustained", ceil["bf16_derived_sustained"])
            basis = ("dense TF32 measured in this run (torch.matmul fp32 8192^3, allow_tf32, back to back) / 3 MMA passes" if tf32
                     else f"{peak_src} bf16_tflops_sustained / 2 / 3")
        else:
            peak, basis = ffma_peak, "148 SM x 128 FMA x 2 x sampled SM clock"
        traffic, tnote = traffic_from_profile(workload, world, mode, nb_outer, pkg)
        roofline = {"bound": "tensor", "kernel": "tc_gemm_pair_kernel / tc_gemm_kernel (trailing update A22 -= L21*U12)" if mode == pkg.MODE_3XTF32 else "sgemm_nn_kernel",
                    "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic, "traffic_note": tnote,
                    "peak_basis": basis, "peak_tf32_measured": tf32["tf32_tflops_sustained"] if tf32 else None, "launches": upd_launches, "avg_launch_ms": upd_s / max(upd_launches, 1) * 1e3,
                    "share_of_step": upd_s / max(fac_s + bs_s + rest_s, 1e-12),
                    "algorithmic_flops_per_step": upd_flop / steps}
        if mode == pkg.MODE_3XTF32:
            roofline["ceilings_tflops"] = ceil
            roofline["frac_of"] = {k: ach / v for k, v in ceil.items()}
            roofline["tf32_mma_tflops_issued_by_kernel"] = 3.0 * ach
            if ach > peak:
                roofline["note"] = ("achieved exceeds the library-GEMM-derived ceiling: cuBLAS TF32 under the power cap is not the silicon "
                                    "limit; frac_of.nominal is the fraction of the nominal dense TF32 rate / 3")
        whole = value * 1e-3
        roofline["whole_step_tflops"] = whole
        roofline["whole_step_frac"] = whole / (peak * world)

    line = None
    cpu = None
    if rank == 0 and with_cpu and world == 1:
        cpu = cpu_baseline_gauss(cores, n if cpu_same_n else None)
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
                "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32 (3xTF32 tensor-core trailing update)" if mode == pkg.MODE_3XTF32 else "f32",
                "data": "synthetic", "impl": "ours",
                "config": {"workload": workload, "n": n, "mode": args.mode, "outer_block": nb_outer, "inner_block": 128,
                           "seed": args.seed,
                           "l2": ("256 MB flush buffer written between timed iterations" if flush_l2 else
                                  "working set (%.1f GiB) exceeds the 126 MB L2; no flush needed" % (n * n * 4 / 2 ** 30)),
                           "step": "restore D2D + computeGauss + solveGauss on the resident system",
                           "parallelism": f"block-cyclic rows x{world}" if world > 1 else "single GPU"},
                "breakdown_ms": {"restore": rest_s / steps * 1e3, "factor": fac_s / steps * 1e3,
                                 "backsub": bs_s / steps * 1e3, "wall_per_step": wall_s / steps * 1e3},
                "factor_only_gflops": flop / (fac_s / steps) * 1e-9,
                "max_rel_err": parity["max_rel_err"], "parity": parity, "graph_replayed_steps": graph_steps,
                "gpu_launches": total_launches, "e2e": e2e, "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks}
        if nvlink:
            line["nvlink"] = nvlink
    g.close()
    hb.close()
    del flush
    return line, parity["ok"]


def traffic_from_profile(workload, world, mode, nb_outer, pkg):
    """DRAM bytes per trailing-update launch from the committed ncu capture -- only when that capture was made at this
    workload, outer block and kernel source (hash of tc_gemm.cu); otherwise null."""
    import hashlib
    tpath = os.path.join(ROOT, "profiles", "r02", "tc_update_traffic.json")
    if world != 1 or mode != pkg.MODE_3XTF32 or not os.path.exists(tpath):
        return None, None
    rec = json.load(open(tpath))
    src = os.path.join(ROOT, "gaussian_elimination-cuda-openmp-mpi-pthreads_b200", "csrc", "tc_gemm.cu")
    sha = hashlib.sha256(open(src, "rb").read()).hexdigest()[:16]
    if rec.get("workload") != workload or rec.get("nb_outer") != nb_outer or rec.get("tc_gemm_sha16") != sha:
        return None, "profiles/r02/tc_update_traffic.json was captured for another configuration or kernel source"
    return rec.get("traffic_bytes_per_launch"), rec.get("note")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="gauss32768", choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="tc", choices=["tc", "ffma"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nb", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the gauss2048 / matmul2048 block of the default run")
    ap.add_argument("--no-tf32-probe", action="store_true")
    ap.add_argument("--env-sweep", default="", help="tuning aid: KEY=v1,v2,... runs the workload once per value of the library's environment switch KEY "
                                                   "(read when the context is created) and prints the list under 'sweep'")
    ap.add_argument("--no-parity-exit", action="store_true", help="tuning aid: report parity but do not exit 3 (timing diagnostics that skip work)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)

    rank = int(os.environ.get("RANK", "0"))
    # stdout carries exactly ONE JSON line: anything libraries print there (NCCL's version banner, ...)
    # is sent to stderr instead, and the real stdout is kept aside for the final line
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    if args.impl == "reference":
        reference_arm(args, rank)
        return

    import torch
    import __graft_entry__ as ge
    pkg = ge.package()
    if not torch.cuda.is_available() or pkg.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device (this framework has no CPU fallback)")
    D = Dist()
    kind, n = WORKLOADS[args.workload]
    mode = pkg.MODE_3XTF32 if args.mode == "tc" else pkg.MODE_FFMA
    peaks, peak_src = measured_peaks()
    cores = os.cpu_count() or 1

    tf32 = None
    if not args.no_tf32_probe and mode == pkg.MODE_3XTF32:
        if rank == 0:
            tf32 = measure_tf32_peak()
        D.barrier()

    if kind == "matmul":
        if rank == 0:
            emit(bench_matmul(args, pkg, n, mode, peaks, peak_src, tf32, cores, args.steps, args.warmup, not args.no_cpu_baseline))
        D.close()
        return

    small = n <= 8192                       # fits in L2 many times over: flush between iterations, CPU baseline at the same N
    if args.env_sweep:
        key, vals = args.env_sweep.split("=", 1)
        out = []
        for v in vals.split(","):
            os.environ[key] = v
            l, ok = bench_gauss(args, pkg, D, args.workload, n, mode, peaks, peak_src, tf32, cores, args.steps, args.warmup,
                                with_e2e=False, with_cpu=False, flush_l2=small)
            if rank == 0:
                out.append({"env": f"{key}={v}", "value": l["value"], "ms_per_step": l["ms_per_step"], "breakdown_ms": l["breakdown_ms"],
                            "parity": {k: l["parity"][k] for k in ("max_rel_err", "scaled_residual", "ok")},
                            "update_tflops_per_launch": l["roofline"]["achieved"] if l["roofline"] else None})
        if rank == 0:
            emit({"metric": METRIC, "unit": UNIT, "n_gpus": D.world, "config": {"workload": args.workload, "n": n, "nb": args.nb}, "sweep": out})
        D.close()
        return
    line, ok = bench_gauss(args, pkg, D, args.workload, n, mode, peaks, peak_src, tf32, cores, args.steps, args.warmup,
                           with_e2e=not args.no_e2e, with_cpu=not args.no_cpu_baseline, flush_l2=small, cpu_same_n=small)
    if rank == 0 and tf32:
        line["tf32_probe"] = tf32
    # configs 1-3 of BASELINE.json ride along on the default line: N = 2048 elimination (latency path) with the reference's
    # sequential and OpenMP programs at the same N, and the 2048^2 matmul with the reference's OpenMP loop and kernel
    if D.world == 1 and args.workload == "gauss32768" and not args.no_secondary:
        sec = {}
        l2, ok2 = bench_gauss(args, pkg, D, "gauss2048", 2048, mode, peaks, peak_src, tf32, cores, 20, 5,
                              with_e2e=True, with_cpu=not args.no_cpu_baseline, flush_l2=True, cpu_same_n=True)
        sec["gauss2048"] = {k: l2[k] for k in ("value", "unit", "ms_per_step", "steps", "warmup", "config", "breakdown_ms", "parity",
                                               "gpu_launches", "graph_replayed_steps", "e2e", "roofline", "cpu_baseline")}
        mm = bench_matmul(args, pkg, 2048, mode, peaks, peak_src, tf32, cores, 20, 5, not args.no_cpu_baseline)
        sec["matmul2048"] = {k: mm[k] for k in ("metric", "value", "unit", "ms_per_step", "steps", "config", "parity", "gpu_launches", "e2e",
                                                "roofline", "reference_kernel_ms_on_this_gpu", "cpu_baseline")}
        line["secondary"] = sec
        ok = ok and ok2 and mm["parity"]["ok"]
    if rank == 0:
        emit(line)
    D.close()
    if not ok and not args.no_parity_exit:
        sys.stderr.write("bench.py: PARITY FAILED (see the line's parity block)\n")
        sys.exit(3)


def bench_matmul(args, pkg, n, mode, peaks, peak_src, tf32, cores, steps, warmup, with_cpu):
    """FP32 2048^2 matmul (configs[2]): single GPU; value = device-pointer entry, e2e = host-buffer entry."""
    import torch
    A, B = pkg.matmul_fill(n)
    dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
    dC = torch.empty_like(dA)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")     # 256 MB > L2
    stream = torch.cuda.current_stream()
    for _ in range(max(warmup, 3)):
        pkg.matmul_cuda_device(dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), n, mode=mode, stream=stream.cuda_stream)
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    evs = []
    for _ in range(steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        pkg.matmul_cuda_device(dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), n, mode=mode, stream=stream.cuda_stream)
        e1.record(stream)
        evs.append((e0, e1))
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms = sum(a.elapsed_time(b) for a, b in evs) / steps
    flop = 2.0 * n ** 3
    value = flop / (ms * 1e-3) * 1e-9
    # parity: element-wise relative 1e-4 against the double-precision product (north_star), on the timed output
    Cg = dC.cpu().numpy()
    C64 = A.astype(np.float64) @ B.astype(np.float64)
    perr = float(np.max(np.abs(Cg - C64) / np.abs(C64)))
    parity = {"max_rel_err": perr, "tol": 1e-4, "ok": bool(perr <= 1e-4), "checked": "C of the timed launches vs the float64 product, element-wise"}
    e2e_steps = min(steps, 5)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        Ch, ks = pkg.matmul_cuda(A, B, mode=mode)
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    ref_ms = None
    try:
        import oracle_lib as ol
        dR = torch.empty_like(dA)
        for _ in range(2):
            ref_ms = ol.ref_matmul().ref_gpu_matmul(dA.data_ptr(), dB.data_ptr(), dR.data_ptr(), n)
    except OSError:
        pass
    cpu = None
    if with_cpu:
        try:                                 # the reference's own OpenMP loop (cuda_matmul.cu:42-57) on the host cores
            import oracle_lib as ol
            t0 = time.perf_counter()
            ol.ref_cpu_matmul(A, B, omp=True)
            tc = time.perf_counter() - t0
            cpu = {"value": flop / tc * 1e-9, "unit": UNIT, "cores": cores, "kind": "reference",
                   "sample": f"omp_matmul at N={n} (the full workload), one run, {tc:.2f} s"}
        except OSError:
            cpu = None
    ceil = tensor_ceilings(peaks, tf32)
    if mode == pkg.MODE_3XTF32:
        peak = ceil.get("tf32_measured_burst", ceil["bf16_derived_burst"])       # a kernel timed alone: the burst figure
        basis = "dense TF32 measured in this run (torch.matmul fp32 8192^3, best of 10) / 3 MMA passes" if tf32 else f"{peak_src} bf16_tflops / 6"
    else:
        peak = 148 * 128 * 2 * ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6 * 1e-12
        basis = "148 SM x 128 FMA x 2 x sampled SM clock"
    line = {"metric": "matmul_gflops", "value": value, "unit": UNIT, "n_gpus": 1, "steps": steps, "warmup": max(warmup, 3),
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32 (3xTF32)" if mode == pkg.MODE_3XTF32 else "f32", "data": "synthetic (reference fill)", "impl": "ours",
            "config": {"workload": f"matmul{n}", "n": n, "mode": args.mode, "l2": "256 MB flush buffer written between timed iterations"},
            "parity": parity,
            "gpu_launches": 3 * steps if mode == pkg.MODE_3XTF32 else steps,
            "e2e": {"value": flop / e2e_s * 1e-9, "unit": UNIT, "h2d_bytes_per_step": 2 * n * n * 4, "d2h_bytes_per_step": n * n * 4,
                    "ms_per_step": e2e_s * 1e3, "note": "matmul_cuda: cudaMalloc + 2 H2D + kernels + D2H + cudaFree, pageable host memory, like cuda_matmul.cu:135-165"},
            "roofline": {"bound": "tensor", "achieved": value * 1e-3, "peak": peak, "unit": "TFLOP/s", "frac": value * 1e-3 / peak,
                         "traffic": None, "peak_basis": basis, "note": "includes the two operand split/transposes",
                         "ceilings_tflops": ceil if mode == pkg.MODE_3XTF32 else None},
            "reference_kernel_ms_on_this_gpu": ref_ms, "cpu_baseline": cpu, "clocks": clocks}
    return line


if __name__ == "__main__":
    main()
