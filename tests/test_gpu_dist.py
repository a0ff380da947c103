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


@pytest.mark.skipif(pkg.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("mode", [pkg.MODE_FFMA, pkg.MODE_3XTF32])
@pytest.mark.parametrize("n,nb", [(1500, 256), (2048, 512), (700, 128)])
def test_two_gpus_match_oracle(tmp_path, n, nb, mode):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n, nb, mode, str(tmp_path)), nprocs=world, join=True)
    xs = []
    for r in range(world):
        errx, errb, same, bytes_rx, w, resid = np.load(tmp_path / f"r{r}.npy")
        assert errx <= 1e-3 and errb <= 1e-3 and resid <= 10.0, (r, errx, errb, resid)
        assert same == 1.0 and w == world
        assert bytes_rx > 0
        xs.append(np.load(tmp_path / f"x{r}.npy"))
    assert np.array_equal(xs[0], xs[1])                      # X is complete and identical on every rank
    with pkg.GaussCuda(n, nb=nb, mode=mode) as g1:
        A, B, _ = pkg.initialize_inputs(n, 6)
        X1 = g1.gauss(A, B)
    assert np.max(np.abs(X1 - xs[0]) / np.abs(X1)) <= 1e-3
