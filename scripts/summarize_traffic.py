#!/usr/bin/env python
"""profiles/r02/tc_update_traffic.json from an ncu CSV of tc_gemm_pair_kernel launches of bench.py --workload gauss32768
(metrics dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum, sm__pipe_tensor_cycles_active...).
The trailing updates (k = outer block) are told from the in-panel rank-128 launches of the same kernel by their duration."""
import collections, csv, hashlib, io, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02", "traffic_pair_gauss32768.csv")
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
lines = open(src).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
per = collections.defaultdict(dict)
for r in csv.DictReader(io.StringIO("\n".join(lines[start:]))):
    per[r["ID"]][r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
P3 = [d for _, d in sorted(per.items(), key=lambda kv: int(kv[0])) if d["gpu__time_duration.sum"] > 0.15e6]     # ns
small = [d for d in per.values() if d["gpu__time_duration.sum"] <= 0.15e6]
n, ld = 32768, 32768 + 32
alg = [8.0 * (n - (K + 1) * nb) * (ld - (K + 1) * nb) for K in range(n // nb - 1)]
# the captured run holds whole solves back to back: match every captured trailing update with its algorithmic bytes
solves = max(1, round(len(P3) / max(1, sum(1 for a in alg if a > 0))))
tm = sum(d["gpu__time_duration.sum"] for d in P3)
per_solve = len(P3) // solves                      # the smallest trailing updates of a solve run on the single-CTA kernel
alg_captured = sum(alg[:per_solve]) * solves
traffic_captured = sum(d["dram__bytes_read.sum"] + d["dram__bytes_write.sum"] for d in P3)
ratio = traffic_captured / alg_captured
rec = {
    "workload": "gauss32768", "nb_outer": nb,
    "tc_gemm_sha16": hashlib.sha256(open(os.path.join(ROOT, "gaussian_elimination-cuda-openmp-mpi-pthreads_b200", "csrc", "tc_gemm.cu"), "rb").read()).hexdigest()[:16],
    "launches_captured": len(P3), "in_panel_launches_of_the_same_kernel_skipped": len(small),
    "solves_captured": solves, "trailing_updates_per_solve_on_this_kernel": per_solve, "trailing_updates_per_solve": len(alg),
    "traffic_over_algorithmic": ratio,
    # bench.py quotes achieved flops per launch averaged over ALL trailing updates of a solve: the same average for the bytes
    "traffic_bytes_per_launch": ratio * sum(alg) / len(alg),
    "captured_traffic_bytes_per_launch": traffic_captured / len(P3), "captured_algorithmic_bytes_per_launch": alg_captured / len(P3),
    "dram_read_bytes_per_launch": sum(d["dram__bytes_read.sum"] for d in P3) / len(P3),
    "dram_write_bytes_per_launch": sum(d["dram__bytes_write.sum"] for d in P3) / len(P3),
    "tensor_pipe_active_pct_time_weighted": sum(d["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"] * d["gpu__time_duration.sum"] for d in P3) / tm,
    "ncu_time_per_launch_ms": tm / len(P3) * 1e-6,
}
first = P3[0]
rec["first_launch"] = {"traffic_bytes": first["dram__bytes_read.sum"] + first["dram__bytes_write.sum"], "algorithmic_bytes": alg[0],
                       "ratio": (first["dram__bytes_read.sum"] + first["dram__bytes_write.sum"]) / alg[0]}
rec["note"] = (f"ncu dram__bytes_read.sum + dram__bytes_write.sum over {len(P3)} trailing-update launches (tc_gemm_pair_kernel, k = {nb}) of bench.py "
               f"--workload gauss32768 ({per_solve} of the {len(alg)} per solve run on this kernel, the {len(alg) - per_solve} smallest on the single-CTA one): "
               f"{ratio:.3f} x the algorithmic bytes (C read + write = 8 m n per launch); traffic_bytes_per_launch = that ratio x the "
               f"full-solve average of {sum(alg) / len(alg):.3e} algorithmic bytes per launch")
json.dump(rec, open(os.path.join(ROOT, "profiles", "r02", "tc_update_traffic.json"), "w"), indent=1)
print(json.dumps(rec, indent=1))
