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


# ---- the critical-path schedule (factor_cp in gauss_cuda.cu), same structure in numpy over gloo ------------------
def _ext_factor(A11):
    """Eliminates the first kb pivots of E = [[A11, I], [I, 0]] (lu_cp.cu): returns L11|U11, inv(L11), inv(U11)."""
    kb = A11.shape[0]
    E = np.zeros((2 * kb, 2 * kb))
    E[:kb, :kb] = A11
    E[:kb, kb:] = np.eye(kb)
    E[kb:, :kb] = np.eye(kb)
    for i in range(kb):
        E[i, i + 1:] /= E[i, i]                              # getPivot :100-111 (the pivot itself stays in L)
        m = E[i + 1:, i].copy()
        E[i + 1:, i + 1:] -= np.outer(m, E[i, i + 1:])       # computeGauss :129-136
    return E[:kb, :kb].copy(), E[:kb, kb:].copy(), E[kb:, :kb].copy()


def _worker_cp(rank, world, port, n, nb, near, seed, out_dir):
    sys.path.insert(0, HERE)
    from _pkg import pkg
    import oracle_lib as ol
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        A, B, Xtrue = ol.initialize_inputs(n, seed)
        npad = pkg.padded_order(n)
        ld = npad + 1
        M = np.zeros((npad, ld))
        M[:n, :n] = A
        M[:n, npad] = B
        for g in range(n, npad):
            M[g, g] = 1.0
        rows = pkg.local_row_map(n, nb, rank, world)
        W = M[rows].copy() if len(rows) else np.zeros((0, ld))
        blocks = pkg.outer_blocks(n, nb)
        nblk = len(blocks)
        lpos = {int(g): i for i, g in enumerate(rows)}
        NB = min(nb, npad)
        inv_u = {}                                           # owner keeps inv(U11) of its blocks for back substitution

        def bcast(arr, shape, src):
            t = torch.from_numpy(np.ascontiguousarray(arr)) if rank == src else torch.empty(shape, dtype=torch.float64)
            dist.broadcast(t, src=src)
            return t.numpy()

        def first_local_below(K):                            # first local row strictly below outer block K
            k_end = blocks[K][0] + blocks[K][1]
            idx = [i for i, g in enumerate(rows) if g >= k_end]
            return idx[0] if idx else len(rows)

        def panel(K):                                        # owner: diagonal region, then the row panel in three column ranges
            k0, kb = blocks[K]
            l = lpos[k0]
            LU, iL, iU = _ext_factor(W[l:l + kb, k0:k0 + kb])
            W[l:l + kb, k0:k0 + kb] = LU
            inv_u[K] = iU
            c1 = k0 + kb
            ut = (iL @ W[l:l + kb, c1:]).T                   # U12^T = A12^T inv(L11)^T, one product (columns c1 .. ld)
            W[l:l + kb, c1:] = ut.T                          # U12 itself stays with the owner for back substitution
            return iU.T.copy(), ut

        crit = ut_near = ut_far = None
        for K in range(nblk):
            k0, kb = blocks[K]
            owner = pkg.block_owner(K, world)
            if K == 0 and owner == rank:
                pend = panel(0)
            if K + 1 >= nblk:
                break
            c1 = k0 + kb
            ccrit = blocks[K + 1][1]
            cn = min(ld, (K + 2 + near) * NB)
            mine = owner == rank
            iut = bcast(pend[0] if mine else None, (kb, kb), owner)                       # crit package, part 1
            uc = bcast(pend[1][:ccrit] if mine else None, (ccrit, kb), owner)             # crit package, part 2: U[K, K+1]^T
            ut_near = bcast(pend[1][:cn - c1] if mine else None, (cn - c1, kb), owner)    # near columns (crit block included)
            ut_far = bcast(pend[1][cn - c1:] if mine else None, (ld - cn, kb), owner) if ld > cn else np.zeros((0, kb))
            l0 = first_local_below(K)
            own_next = pkg.block_owner(K + 1, world) == rank
            l1 = l0 + (ccrit if own_next else 0)
            # L21 = A21 inv(U11); the columns of block K+1 from the crit package
            W[l0:, k0:c1] = W[l0:, k0:c1] @ iut.T
            L21 = W[l0:, k0:c1]
            W[l0:, c1:c1 + ccrit] -= L21 @ uc.T
            ub = np.vstack([ut_near, ut_far])                # rows = columns c1 .. ld of the matrix
            assert np.array_equal(ub[:ccrit], uc)
            if own_next:                                     # the rest of block K+1's rows, then its panel
                W[l0:l1, c1 + ccrit:] -= W[l0:l1, k0:c1] @ ub[ccrit:].T
                pend = panel(K + 1)
            W[l1:, c1 + ccrit:] -= W[l1:, k0:c1] @ ub[ccrit:].T
        # back substitution with the stored inverses, last outer block first
        x = np.zeros(npad)
        for K in range(nblk - 1, -1, -1):
            k0, kb = blocks[K]
            owner = pkg.block_owner(K, world)
            xk = bcast(inv_u[K] @ W[lpos[k0]:lpos[k0] + kb, npad] if owner == rank else None, (kb,), owner)
            x[k0:k0 + kb] = xk
            above = [i for i, g in enumerate(rows) if g < k0]
            if above:
                W[:above[-1] + 1, npad] -= W[:above[-1] + 1, k0:k0 + kb] @ xk
        Xref, _, _ = ol.gauss(A, B)
        err = float(np.max(np.abs(x[:n] - Xref) / np.abs(Xref)))
        np.save(os.path.join(out_dir, f"err{rank}.npy"), np.array([err, float(np.abs(x[n:]).max() if npad > n else 0.0)]))
    finally:
        dist.destroy_process_group()


# (4, 300, 128): three outer blocks on four ranks -- rank 3 owns no rows and still takes part in every broadcast
@pytest.mark.parametrize("world,n,nb,near", [(2, 300, 128, 1), (2, 700, 256, 3), (3, 900, 128, 2), (4, 300, 128, 3)])
def test_critical_path_schedule_matches_oracle(tmp_path, world, n, nb, near):
    port = _free_port()
    mp.spawn(_worker_cp, args=(world, port, n, nb, near, 4, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        err, padmax = np.load(tmp_path / f"err{r}.npy")
        assert err < 1e-9, (r, err)
        assert padmax == 0.0
