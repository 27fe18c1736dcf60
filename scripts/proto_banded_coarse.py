"""CPU design check for the NEXT coarse inverse: the Galerkin matrix of the
coordinate-bin coarse space is a sparse grid operator, and the dense inverse the iteration applies can be obtained
from a BANDED factorisation instead of a dense Gauss-Jordan.

What it does, on the reference's matrix (oracle assembly) with the same Jacobi + coordinate-bin coarse space as
csrc/pdn_coarse.cuh:
  1. A_c = P^T A P: stored entries per row, bandwidth in the bin numbering the CUDA code uses (mesh-major) and after
     reverse Cuthill-McKee;
  2. banded LU without pivoting in the RCM order + one banded solve per column of the identity = the dense inverse;
     compared entry-wise with numpy's dense inverse (and after rounding to fp32, as the iteration applies it);
  3. operation counts: dense Gauss-Jordan 2 n^3 against banded n b^2 (factor) + 2 n^2 (2 b + 1) (all columns), and the
     length of the serial chain (block steps of the current kernel against rows of a banded factor).

    python scripts/proto_banded_coarse.py [layers: 2|8] [grid_side] [bins_per_side]
"""
import os
import sys
import time

import numpy as np
import scipy.linalg
import scipy.sparse as sp
import scipy.sparse.csgraph as csgraph
import scipy.sparse.linalg as spla

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle  # noqa: E402
from padne_b200 import synth  # noqa: E402
from padne_b200.stamps import build_stamps  # noqa: E402

layers = int(sys.argv[1]) if len(sys.argv) > 1 else 2
side = int(sys.argv[2]) if len(sys.argv) > 2 else (224 if layers == 2 else 110)
bins = int(sys.argv[3]) if len(sys.argv) > 3 else (27 if layers == 2 else 14)

board = (synth.board_2layer(side, n_via=16, size_mm=100.0, seed=0) if layers == 2
         else synth.board_8layer(side, n_via=8, size_mm=300.0, seed=0))
fs = board.to_flat()
st = build_stamps(fs)
indptr, indices, data = oracle.assemble(fs, st[:4])
L = sp.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
n = fs.n_nodes
vs = np.flatnonzero(fs.e_kind == 2)
fixed = np.unique(np.concatenate([fs.e_node[vs, 0], fs.e_node[vs, 1], [fs.i_gnd]]))
free = np.setdiff1d(np.arange(n), fixed)
A = (-L[:n, :n]).tocsr()[free][:, free].tocsr()
xy = fs.xy[np.minimum(free, fs.n_vertices - 1)]
mesh_of = np.searchsorted(fs.v_off, np.minimum(free, fs.n_vertices - 1), side="right") - 1
span = xy.max(0) - xy.min(0) + 1e-9
cell = np.minimum((bins * (xy - xy.min(0)) / span).astype(int), bins - 1)
raw = mesh_of * bins * bins + cell[:, 1] * bins + cell[:, 0]                    # mesh-major, then rows of bins
raw[free >= fs.n_vertices] = mesh_of.max() * bins * bins + bins * bins + np.arange((free >= fs.n_vertices).sum())  # own aggregates
agg = np.unique(raw, return_inverse=True)[1]
nc = int(agg.max()) + 1
P = sp.csr_matrix((np.ones(len(free)), (np.arange(len(free)), agg)), shape=(len(free), nc))
Ac = (P.T @ A @ P).tocsr()
Ac.sum_duplicates()
print(f"{layers}-layer board, grid {side}: {len(free)} rows, {bins} x {bins} bins per layer -> n_c = {nc}")
per_row = np.diff(Ac.indptr)
print(f"A_c: {Ac.nnz} stored entries = {Ac.nnz / nc:.1f} per row (max {per_row.max()}), {Ac.nnz / nc ** 2:.2%} of the dense matrix")


def bandwidth(M):
    coo = M.tocoo()
    return int(np.abs(coo.row - coo.col).max())


perm = csgraph.reverse_cuthill_mckee(Ac, symmetric_mode=True)
Ar = Ac[perm][:, perm].tocsr()
b0, b = bandwidth(Ac), bandwidth(Ar)
print(f"half-bandwidth: {b0} in the mesh-major bin numbering, {b} after reverse Cuthill-McKee")

# banded LU + one solve per column = the dense inverse (scipy's solve_banded pivots inside the band; the CUDA design is
# no pivoting, like today's Gauss-Jordan -- A_c is an M-matrix; checked separately below with an explicit no-pivot LU)
ab = np.zeros((2 * b + 1, nc))
coo = Ar.tocoo()
ab[b + coo.row - coo.col, coo.col] = coo.data
t0 = time.time()
inv_r = scipy.linalg.solve_banded((b, b), ab, np.eye(nc))
t_band = time.time() - t0
inv_banded = np.empty_like(inv_r)
inv_banded[np.ix_(perm, perm)] = inv_r
t0 = time.time()
inv_dense = np.linalg.inv(Ac.toarray())
t_dense = time.time() - t0
scale = np.abs(inv_dense).max()
print(f"inverse from the banded factor vs dense inverse: max |diff| / max |entry| = {np.abs(inv_banded - inv_dense).max() / scale:.2e}; "
      f"after rounding both to fp32: {np.abs(inv_banded.astype(np.float32) - inv_dense.astype(np.float32)).max() / scale:.2e}")
print(f"(CPU, LAPACK: banded {t_band * 1e3:.0f} ms, dense {t_dense * 1e3:.0f} ms)")

# explicit no-pivot banded LU (what a kernel would run): growth and the smallest pivot
Bd = Ar.toarray()
piv_min = np.inf
for k in range(nc - 1):
    hi = min(nc, k + b + 1)
    piv_min = min(piv_min, Bd[k, k])
    l = Bd[k + 1:hi, k] / Bd[k, k]
    Bd[k + 1:hi, k + 1:hi] -= np.outer(l, Bd[k, k + 1:hi])
    Bd[k + 1:hi, k] = l
piv_min = min(piv_min, Bd[-1, -1])
print(f"no-pivot banded LU: smallest pivot {piv_min:.3e} (diagonal of A_c: {Ac.diagonal().min():.3e} .. {Ac.diagonal().max():.3e}), "
      f"max |factor entry| {np.abs(Bd).max():.3e}")

flops_gj = 2.0 * nc ** 3
flops_band = nc * b * b * 2.0 + 2.0 * nc * nc * (2 * b + 1)
print(f"flops: dense Gauss-Jordan {flops_gj / 1e9:.2f} G, banded factor + {nc} column solves {flops_band / 1e9:.3f} G ({flops_gj / flops_band:.0f}x fewer)")
print(f"serial chain: {-(-nc // 12)} block steps of ~10.5 us today (1-D row blocks, profiles/r2_coarse_inverse.txt) against "
      f"{nc} pivot rows of a {b}-wide band for the factor; the {nc} column solves are independent (one warp each)")

# does the preconditioner care?  PCG iterations with the exact inverse and with the banded one (both applied from fp32)
dinv = 1.0 / A.diagonal()
rng = np.random.default_rng(0)
rhs = A @ rng.standard_normal(A.shape[0])


def pcg(inv_c):
    inv32 = inv_c.astype(np.float32).astype(np.float64)
    x = np.zeros_like(rhs); r = rhs.copy()
    z = dinv * r + P @ (inv32 @ (P.T @ r)); p = z.copy(); rz = r @ z
    for it in range(1, 2000):
        w = A @ p; alpha = rz / (p @ w); x += alpha * p; r -= alpha * w
        if np.linalg.norm(r) <= 1e-10 * np.linalg.norm(rhs):
            return it
        z = dinv * r + P @ (inv32 @ (P.T @ r)); rz_new = r @ z; p = z + (rz_new / rz) * p; rz = rz_new
    return -1


print(f"PCG to 1e-10 with the two inverses: {pcg(inv_dense)} and {pcg(inv_banded)} iterations")
