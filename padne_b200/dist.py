"""One process per GPU: wiring a :class:`padne_b200.capi.Context` into a ``torch.distributed`` job.

``torch.distributed`` (NCCL or gloo) is only the transport for the 128-byte NCCL id and for host-side
bookkeeping; the data-path collectives (halo exchange, dot-product all-reduce) are issued by
``libpadne_b200.so`` itself on its own stream (``csrc/pdn_multi.cu``), captured in the same CUDA graph
as the Krylov kernels.

The partition used by the solver (``csrc/pdn_solve.cu``) is restated here as :func:`row_partition` so
host code and tests can reason about it without a GPU.
"""
from __future__ import annotations

import numpy as np

SELL_WIN = 1024     # rows per sort window (csrc/pdn_sell.cuh); ranks own whole windows


def row_partition(n: int, world: int) -> np.ndarray:
    """``begin[k]`` = first row (permuted numbering) of rank ``k``; ``begin[world] = n``."""
    n_win = (n + SELL_WIN - 1) // SELL_WIN
    begin = np.minimum(n, (np.arange(world + 1, dtype=np.int64) * n_win // world) * SELL_WIN)
    begin[world] = n
    return begin


def owner_of(begin: np.ndarray, rows) -> np.ndarray:
    """Rank owning each row."""
    return np.searchsorted(begin, np.asarray(rows), side="right") - 1


def halo_lists(indptr, indices, begin, me: int):
    """Send / receive index lists of rank ``me`` for a CSR operator whose rows are split at ``begin``.

    Mirrors ``build_halo_plan``: rank ``k`` needs column ``c`` from its owner when one of ``k``'s rows
    stores an entry in ``c`` and ``c`` is not ``k``'s.  Returns ``(send, recv)``: dicts peer ->
    ascending int64 index array."""
    world = len(begin) - 1
    indptr = np.asarray(indptr)
    indices = np.asarray(indices)
    rows = np.repeat(np.arange(len(indptr) - 1), np.diff(indptr))
    r_own = owner_of(begin, rows)
    c_own = owner_of(begin, indices)
    cross = r_own != c_own
    send, recv = {}, {}
    for k in range(world):
        if k == me:
            continue
        s = np.unique(indices[cross & (c_own == me) & (r_own == k)])
        r = np.unique(indices[cross & (r_own == me) & (c_own == k)])
        if s.size:
            send[k] = s.astype(np.int64)
        if r.size:
            recv[k] = r.astype(np.int64)
    return send, recv


def broadcast_unique_id(make_id, rank: int, group=None) -> bytes:
    """Rank 0 calls ``make_id()``; everybody returns the same 128 bytes (works on gloo and nccl)."""
    import torch
    import torch.distributed as dist
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    t = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        t.copy_(torch.frombuffer(bytearray(make_id()), dtype=torch.uint8))
    dist.broadcast(t, src=0, group=group)
    return bytes(t.cpu().numpy().tobytes())


def init_comm(ctx, group=None) -> tuple[int, int]:
    """Give ``ctx`` a communicator spanning the ``torch.distributed`` job.  Returns ``(rank, world)``."""
    import torch.distributed as dist
    if not dist.is_initialized():
        return 0, 1
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if world == 1:
        return 0, 1
    uid = broadcast_unique_id(type(ctx).comm_unique_id, rank, group)
    ctx.comm_init(rank, world, uid)
    return rank, world


def local_level_aggregates(begin: np.ndarray, me: int, raw_bin: np.ndarray) -> tuple[np.ndarray, int]:
    """Aggregate ids of the rank-local second coarse level (``build_coarse(..., local=true)`` in
    ``csrc/pdn_solve.cu``), restated for host code and tests: every row owned by rank ``me`` keeps its
    (fine) coordinate bin ``raw_bin[row]`` (negative = none: a row the constraints eliminated), rows of
    other ranks get -1, and the surviving bins are renumbered compactly in ascending bin order.  The
    level ``P_k (P_k^T A P_k)^-1 P_k^T`` built from these ids only couples rows of rank ``me``, which is why it
    needs no exchange.  Returns (agg [n], number of aggregates)."""
    raw_bin = np.asarray(raw_bin, np.int64)
    rows = np.arange(raw_bin.shape[0])
    mine = (rows >= begin[me]) & (rows < begin[me + 1]) & (raw_bin >= 0)
    agg = np.full(raw_bin.shape[0], -1, np.int64)
    used = np.unique(raw_bin[mine])
    agg[mine] = np.searchsorted(used, raw_bin[mine])
    return agg, int(used.shape[0])
