"""Multi-rank host logic on the CPU (gloo, world size 2): the row partition, the halo lists and a
partitioned Jacobi-PCG restatement of what ``csrc/pdn_solve.cu`` + ``csrc/pdn_multi.cu`` do per
iteration (halo exchange of the SpMV input, all-reduce of the fused dot products).  The GPU path itself
is checked on real B200s by ``scripts/dist_check.py``."""
from __future__ import annotations

import os
import socket

import numpy as np
import pytest
import scipy.sparse
import scipy.sparse.linalg

from conftest import golden_laplace_files
from padne_b200 import dist as pdist


def _spd_system():
    d = np.load([p for p in golden_laplace_files() if "holes_16x14" in p][0])
    n = len(d["indptr"]) - 1
    L = scipy.sparse.csr_matrix((d["data"], d["indices"], d["indptr"]), shape=(n, n))
    A = (-L + scipy.sparse.diags(np.where(np.arange(n) % 37 == 0, 50.0, 0.0))).tocsr()   # a few grounded nodes
    A.sort_indices()
    b = np.random.default_rng(5).standard_normal(n)
    return A, b


def test_row_partition_covers_all_rows():
    for n in (1, 1023, 1024, 1025, 5000, 997921):
        for world in (1, 2, 3, 8):
            begin = pdist.row_partition(n, world)
            assert begin[0] == 0 and begin[-1] == n and np.all(np.diff(begin) >= 0)
            assert np.all(begin[1:-1] % pdist.SELL_WIN == 0)
            if n >= world * pdist.SELL_WIN * 4:
                assert np.diff(begin).max() - np.diff(begin).min() <= pdist.SELL_WIN
    assert list(pdist.owner_of(np.array([0, 1024, 3000]), [0, 1023, 1024, 2999])) == [0, 0, 1, 1]


def test_halo_lists_are_consistent_between_ranks():
    A, _ = _spd_system()
    begin = np.array([0, 64, 150, A.shape[0]])          # ragged 3-way split
    plans = [pdist.halo_lists(A.indptr, A.indices, begin, me) for me in range(3)]
    for me in range(3):
        send, recv = plans[me]
        for peer, idx in send.items():
            assert np.array_equal(idx, plans[peer][1][me])            # my send list == the peer's recv list
            assert np.all(pdist.owner_of(begin, idx) == me)
        for peer, idx in recv.items():
            assert np.all(pdist.owner_of(begin, idx) == peer)
    # every off-rank column a rank touches is received
    for me in range(3):
        rows = slice(begin[me], begin[me + 1])
        cols = np.unique(A[rows].indices)
        foreign = cols[pdist.owner_of(begin, cols) != me]
        got = np.concatenate(list(plans[me][1].values())) if plans[me][1] else np.zeros(0, np.int64)
        assert np.array_equal(np.sort(got), foreign)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _pcg_rank(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        uid = pdist.broadcast_unique_id(lambda: bytes(range(128)), rank)      # the transport of the NCCL id
        assert uid == bytes(range(128))
        A, b = _spd_system()
        n = A.shape[0]
        begin = np.array([0, 96, n])
        lo, hi = begin[rank], begin[rank + 1]
        send, recv = pdist.halo_lists(A.indptr, A.indices, begin, rank)
        Aown = A[lo:hi]
        dinv = 1.0 / A.diagonal()

        def exchange(vec):          # owner pushes its interface entries into the peer's full-length copy
            reqs = []
            for peer, idx in send.items():
                reqs.append(dist.isend(torch.from_numpy(vec[idx].copy()), peer))
            for peer, idx in recv.items():
                buf = torch.empty(len(idx), dtype=torch.float64)
                dist.recv(buf, peer)
                vec[idx] = buf.numpy()
            for r in reqs:
                r.wait()

        def allreduce(*vals):
            t = torch.tensor(vals, dtype=torch.float64)
            dist.all_reduce(t)
            return t.tolist()

        # Chronopoulos-Gear PCG, exactly the kernel sequence of enqueue_pcg_iteration
        x = np.zeros(n); r = b.copy(); z = np.zeros(n); p = np.zeros(n); s = np.zeros(n); w = np.zeros(n)
        z[lo:hi] = dinv[lo:hi] * r[lo:hi]
        bnorm2 = float(b @ b)
        gamma_old = alpha = 1.0
        first, iters = True, 0
        while True:
            exchange(z)
            w[lo:hi] = Aown @ z
            gamma, delta, rr = allreduce(r[lo:hi] @ z[lo:hi], z[lo:hi] @ w[lo:hi], r[lo:hi] @ r[lo:hi])
            if rr <= 1e-20 * bnorm2 or iters >= 2000:
                break
            beta = 0.0 if first else gamma / gamma_old
            alpha = gamma / delta if first else gamma / (delta - beta * gamma / alpha)
            first, gamma_old, iters = False, gamma, iters + 1
            p[lo:hi] = z[lo:hi] + beta * p[lo:hi]
            s[lo:hi] = w[lo:hi] + beta * s[lo:hi]
            x[lo:hi] += alpha * p[lo:hi]
            r[lo:hi] -= alpha * s[lo:hi]
            z[lo:hi] = dinv[lo:hi] * r[lo:hi]
        # ragged slices: broadcast each owner's slice (pdn_comm_allgatherv)
        for k in range(world):
            t = torch.from_numpy(x[begin[k]:begin[k + 1]].copy())
            dist.broadcast(t, src=k)
            x[begin[k]:begin[k + 1]] = t.numpy()
        out.put((rank, iters, x))
    finally:
        dist.destroy_process_group()


def test_partitioned_pcg_matches_single_rank_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_pcg_rank, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = sorted(out.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    A, b = _spd_system()
    ref = scipy.sparse.linalg.spsolve(A.tocsc(), b)
    (_, it0, x0), (_, it1, x1) = results
    assert it0 == it1 and 0 < it0 < 2000
    assert np.array_equal(x0, x1)                                   # identical bits on both ranks
    assert np.max(np.abs(x0 - ref)) <= 1e-8 * np.max(np.abs(ref))


def test_rank_local_coarse_level_restated():
    # The preconditioner of the partitioned PCG from 4 ranks on: Jacobi + a coarse global level
    # + one block-diagonal level per rank with finer bins.  Host restatement on stacked grid Laplacians:
    # it stays symmetric positive definite, needs no coupling between ranks, converges to the same
    # solution and in fewer iterations than the global level alone.
    n_side, layers, world = 64, 4, 4
    e = np.ones(n_side)
    T = scipy.sparse.diags([-e[:-1], 2 * e, -e[:-1]], [-1, 0, 1]).tolil()
    T[0, 0] = T[n_side - 1, n_side - 1] = 1.0
    T = T.tocsr()
    eye = scipy.sparse.identity(n_side)
    plane = scipy.sparse.kron(eye, T) + scipy.sparse.kron(T, eye)
    n1, n = n_side * n_side, layers * n_side * n_side
    A = scipy.sparse.block_diag([plane] * layers).tolil()
    rng = np.random.default_rng(3)
    for layer in range(layers - 1):                        # vias
        for i in rng.integers(0, n1, 6):
            a, b = layer * n1 + i, (layer + 1) * n1 + i
            A[a, a] += 40.0; A[b, b] += 40.0; A[a, b] -= 40.0; A[b, a] -= 40.0
    A[5, 5] += 1e3                                          # ground
    A = A.tocsr()
    rhs = np.zeros(n)
    rhs[n - 7] = 1.0
    d = A.diagonal()
    ij = np.stack(np.meshgrid(np.arange(n_side), np.arange(n_side), indexing="ij"), -1).reshape(-1, 2)

    def bins(per_side):
        h = -(-n_side // per_side)
        one = (ij[:, 0] // h) * per_side + ij[:, 1] // h
        return np.concatenate([one + layer * per_side * per_side for layer in range(layers)])

    def level(agg, nc):
        rows = np.flatnonzero(agg >= 0)
        P = scipy.sparse.csr_matrix((np.ones(rows.shape[0]), (rows, agg[rows])), shape=(n, nc))
        Ac = (P.T @ A @ P).toarray()
        assert np.allclose(Ac, Ac.T) and np.all(np.linalg.eigvalsh(Ac) > 0)
        return P, np.linalg.inv(Ac)

    begin = pdist.row_partition(n, world)
    assert begin[-1] == n and np.all(np.diff(begin) > 0)
    g_agg = bins(2)
    levels_global = [level(g_agg, int(g_agg.max()) + 1)]
    levels_local = []
    for me in range(world):
        agg, nc = pdist.local_level_aggregates(begin, me, bins(8))
        assert np.all(agg[:begin[me]] == -1) and np.all(agg[begin[me + 1]:] == -1) and nc == agg.max() + 1
        P, inv = level(agg, nc)
        assert P[:begin[me]].nnz == 0 and P[begin[me + 1]:].nnz == 0       # only this rank's rows: no exchange
        levels_local.append((P, inv))

    def solve(levels):
        count = [0]

        def apply(r):
            z = r / d
            for P, inv in levels:
                z = z + P @ (inv @ (P.T @ r))
            return z

        M = scipy.sparse.linalg.LinearOperator((n, n), matvec=apply)
        x, info = scipy.sparse.linalg.cg(A, rhs, rtol=1e-10, maxiter=5000, M=M,
                                         callback=lambda _: count.__setitem__(0, count[0] + 1))
        assert info == 0
        return x, count[0]

    x_g, it_g = solve(levels_global)
    x_l, it_l = solve(levels_global + levels_local)
    assert it_l < 0.8 * it_g
    assert np.max(np.abs(x_l - x_g)) <= 1e-7 * np.max(np.abs(x_g))
