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
