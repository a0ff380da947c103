"""world_size-2 (and 3) CPU tests of the multi-GPU host logic over gloo.

The GPU library shards [A|b] as block-cyclic row panels and broadcasts each factored panel
(gauss_cuda.cu, factor()/backsub()).  Here the same schedule -- same ownership functions from the
package, same broadcast points -- is executed with numpy arithmetic in double and torch.distributed
(gloo) broadcasts, and its result must equal the oracle's.  This pins the ownership/ordering logic
without a GPU; the arithmetic kernels themselves are covered by the -m gpu tests.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _panel_factor(P, k0):
    """Crout elimination restricted to the panel rows (getPivot :100-111 + computeGauss :129-136)."""
    kb = P.shape[0]
    for i in range(kb):
        piv = P[i, k0 + i]
        P[i, k0 + i + 1:] /= piv
        for j in range(i + 1, kb):
            m = P[j, k0 + i]
            P[j, k0 + i + 1:] -= m * P[i, k0 + i + 1:]


def _worker(rank, world, port, n, nb, seed, out_dir):
    sys.path.insert(0, HERE)
    from _pkg import pkg
    import oracle_lib as ol
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        # the id courier used by init_distributed: rank 0 draws it, everyone must end up with the same 128 bytes
        payload = [pkg.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(payload, src=0)
        assert isinstance(payload[0], bytes) and len(payload[0]) == 128

        A, B, Xtrue = ol.initialize_inputs(n, seed)
        npad = pkg.padded_order(n)
        M = np.zeros((npad, npad + 1))
        M[:n, :n] = A
        M[:n, npad] = B
        for g in range(n, npad):
            M[g, g] = 1.0
        rows = pkg.local_row_map(n, nb, rank, world)
        W = M[rows].copy()                                  # this rank's rows only
        blocks = pkg.outer_blocks(n, nb)
        lpos = {int(g): i for i, g in enumerate(rows)}
        for K, (k0, kb) in enumerate(blocks):
            owner = pkg.block_owner(K, world)
            if owner == rank:
                l0 = lpos[k0]
                _panel_factor(W[l0:l0 + kb], k0)
                panel = torch.from_numpy(W[l0:l0 + kb, k0:].copy())
            else:
                panel = torch.empty((kb, npad + 1 - k0), dtype=torch.float64)
            dist.broadcast(panel, src=owner)                # gauss_mpi :137 -> one broadcast per outer block
            Pn = panel.numpy()
            below = [i for i, g in enumerate(rows) if g >= k0 + kb]
            if below:
                lo = below[0]
                assert below == list(range(lo, len(rows)))   # rows below a panel are a contiguous local suffix
                A21 = W[lo:, k0:k0 + kb]
                for j in range(kb):                          # L21 = A21 * U11^-1 (unit upper)
                    A21[:, j] -= A21[:, :j] @ Pn[:j, j]
                W[lo:, k0 + kb:] -= A21 @ Pn[:, kb:]
        # back substitution, last outer block first
        x = np.zeros(npad)
        for K in range(len(blocks) - 1, -1, -1):
            k0, kb = blocks[K]
            owner = pkg.block_owner(K, world)
            xk = torch.empty(kb, dtype=torch.float64)
            if owner == rank:
                l0 = lpos[k0]
                for i in range(kb - 1, -1, -1):
                    W[l0 + i, npad] -= W[l0 + i, k0 + i + 1:k0 + kb] @ W[l0 + i + 1:l0 + kb, npad]
                xk = torch.from_numpy(W[l0:l0 + kb, npad].copy())
            dist.broadcast(xk, src=owner)
            x[k0:k0 + kb] = xk.numpy()
            above = [i for i, g in enumerate(rows) if g < k0]
            if above:
                assert above == list(range(0, above[-1] + 1))
                W[:above[-1] + 1, npad] -= W[:above[-1] + 1, k0:k0 + kb] @ x[k0:k0 + kb]
        Xref, _, _ = ol.gauss(A, B)
        err = float(np.max(np.abs(x[:n] - Xref) / np.abs(Xref)))
        np.save(os.path.join(out_dir, f"err{rank}.npy"), np.array([err, float(np.abs(x[n:]).max() if npad > n else 0.0)]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,nb", [(2, 300, 128), (2, 700, 256), (3, 520, 128)])
def test_block_cyclic_schedule_matches_oracle(tmp_path, world, n, nb):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, nb, 4, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        err, padmax = np.load(tmp_path / f"err{r}.npy")
        assert err < 1e-10, (r, err)
        assert padmax == 0.0
