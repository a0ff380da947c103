"""Multi-GPU parity (needs >= 2 GPUs; skipped on a 1-GPU box): every rank's X from the block-cyclic
NCCL path must match the single-GPU result and the oracle within the north_star tolerance."""
import os
import socket
import sys

import numpy as np
import pytest

from _pkg import pkg

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worlds():
    have = pkg.device_count()
    return [w for w in (2, 4, 8) if w <= have] or [2]


def _worker(rank, world, port, n, nb, mode, out_dir):
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist
    from _pkg import pkg as P
    import oracle_lib as ol
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        A, B, Xtrue = P.initialize_inputs(n, 6)
        g = P.init_distributed(n, nb=nb, mode=mode, device=rank)
        X, Bafter = g.gauss(A, B, want_b_after=True)
        st = g.stats
        g.upload(A, B)
        g.solve_resident()
        X2 = g.download()
        g.close()
        Xref, Bref, _ = ol.gauss(A, B, nthreads=4)
        np.save(os.path.join(out_dir, f"r{rank}.npy"),
                np.array([ol.max_rel_error(X, Xref), ol.max_rel_error(Bafter, Bref), float(np.array_equal(X, X2)),
                          st.bcast_bytes, st.world, ol.scaled_residual(A, B, X)]))
        np.save(os.path.join(out_dir, f"x{rank}.npy"), X)
    finally:
        dist.destroy_process_group()


# (n, nb): full and ragged outer blocks; (300, 128) has 3 outer blocks, so from world = 4 some ranks own NO rows and
# must still take part in every collective; (1100, 256) is 5 blocks (ragged over 4 ranks); (640, 0) = automatic block
@pytest.mark.skipif(pkg.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("world", _worlds())
@pytest.mark.parametrize("mode", [pkg.MODE_FFMA, pkg.MODE_3XTF32])
@pytest.mark.parametrize("n,nb", [(1500, 256), (2048, 512), (700, 128), (300, 128), (1100, 256), (100, 128)])
def test_ranks_match_oracle(tmp_path, n, nb, mode, world):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(world, _free_port(), n, nb, mode, str(tmp_path)), nprocs=world, join=True)
    xs = []
    for r in range(world):
        errx, errb, same, bytes_rx, w, resid = np.load(tmp_path / f"r{r}.npy")
        assert errx <= 1e-3 and errb <= 1e-3 and resid <= 10.0, (r, errx, errb, resid)
        assert same == 1.0 and w == world
        xs.append(np.load(tmp_path / f"x{r}.npy"))
    assert all(np.array_equal(xs[0], x) for x in xs[1:])     # X is complete and identical on every rank
    with pkg.GaussCuda(n, nb=nb, mode=mode) as g1:
        A, B, _ = pkg.initialize_inputs(n, 6)
        X1 = g1.gauss(A, B)
    assert np.max(np.abs(X1 - xs[0]) / np.abs(X1)) <= 1e-3


@pytest.mark.skipif(pkg.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("ngpus", _worlds())
@pytest.mark.parametrize("mode", [pkg.MODE_FFMA, pkg.MODE_3XTF32])
@pytest.mark.parametrize("n", [2048, 700])
def test_single_process_group_matches_oracle(n, mode, ngpus):
    # gauss_cuda_init(ctx, n, ngpus > 1, ...): ONE process, one host thread of the caller, G devices (the reference's
    # multi-worker run is one program too, OpenMP_and_MPI/README.txt:23).  Same tolerance as the one-process-per-GPU form.
    import oracle_lib as ol
    A, B, Xtrue = pkg.initialize_inputs(n, 9)
    with pkg.GaussCuda(n, mode=mode, ngpus=ngpus) as g:
        X, Bafter = g.gauss(A, B, want_b_after=True)
        assert g.stats.world == ngpus
        g.upload(A, B)
        g.solve_resident()
        X2 = g.download()
        g.upload(A, B)
        g.computeGauss()
        g.solveGauss()
        X3 = g.download()
    Xref, Bref, _ = ol.gauss(A, B, nthreads=4)
    assert ol.max_rel_error(X, Xref) <= 1e-3 and ol.max_rel_error(Bafter, Bref) <= 1e-3
    assert ol.scaled_residual(A, B, X) <= 10.0
    assert np.array_equal(X, X2) and np.array_equal(X, X3)


@pytest.mark.skipif(pkg.device_count() < 2, reason="needs 2 GPUs")
def test_harness_drives_two_gpus_from_one_process():
    # harness/gauss_main -s N -g 2 -v : the reference program's output (gauss_internal_input.c:45-48, :225-228) from 2 GPUs
    import re
    import subprocess
    exe = os.path.join(os.path.dirname(HERE), "harness", "gauss_main")
    if not os.path.exists(exe):
        pytest.skip("harness not built")
    n = 4096
    out = subprocess.run([exe, "-s", str(n), "-g", "2", "-v"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, (out.stdout[-500:], out.stderr)
    rows = [l.split() for l in out.stdout.splitlines() if re.match(r"^-?\d+\.\d{5} -?\d+\.\d{5}$", l.strip())]
    assert len(rows) == n
    assert rows[0] == ["0.00000", "-0.50000"] and rows[-1] == ["0.50000", "0.50000"]
    assert all(r == ["0.50000", "0.00000"] for r in rows[1:-1])
    out = subprocess.run([exe, "-g", "2", "3000", "5"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, (out.stdout[-500:], out.stderr)
