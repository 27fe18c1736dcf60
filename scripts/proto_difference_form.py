"""Design check for the next SpMV step, numpy/scipy on the CPU.

The Krylov operator's rows sum to zero except next to eliminated constraint nodes, so the SpMV can skip the stored
diagonal:  (A x)_i = sum_{j != i} a_ij (x_j - x_i) + s_i x_i  with  s_i = sum_j a_ij  (zero for all but a few rows,
which the override path already handles).  Checked here on the reference's matrix (oracle assembly) of a 2-layer
board: how many rows have s_i != 0, how far the two forms differ in one product, and that Jacobi-PCG to 1e-10 takes
the same iterations and reaches the same TRUE residual (computed with the stored diagonal) with either form.

    python scripts/proto_difference_form.py [grid_side]
"""
import os
import sys

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle  # noqa: E402
from padne_b200 import synth  # noqa: E402
from padne_b200.stamps import build_stamps  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 160
board = synth.board_2layer(side, n_via=4, size_mm=100.0, seed=0)
fs = board.to_flat()
st = build_stamps(fs)
indptr, indices, data = oracle.assemble(fs, st[:4])
L = sp.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
v_ref = spla.spsolve(L.tocsc(), st[4])
n = fs.n_nodes
vs = np.flatnonzero(fs.e_kind == 2)
fixed = np.unique(np.concatenate([fs.e_node[vs, 0], fs.e_node[vs, 1], [fs.i_gnd]]))
free = np.setdiff1d(np.arange(n), fixed)
A_full = (-L[:n, :n]).tocsr()
A = A_full[free][:, free].tocsr()
b = -st[4][:n][free] - A_full[free][:, fixed] @ v_ref[fixed]

off = (A - sp.diags(A.diagonal())).tocsr()
off_sum = np.asarray(off.sum(1)).ravel()
row_sum = np.asarray(A.sum(1)).ravel()
row_sum[np.abs(row_sum) <= 1e-12 * A.diagonal()] = 0.0          # exact zeros up to the rounding of the stored diagonal
print(f"rows {A.shape[0]}, rows with a non-zero row sum: {np.count_nonzero(row_sum)}")


def mul_stored(x):
    return A @ x


def mul_difference(x):
    return off @ x - off_sum * x + row_sum * x


x = np.random.default_rng(0).normal(size=A.shape[0])
d = np.linalg.norm(mul_stored(x) - mul_difference(x)) / np.linalg.norm(mul_stored(x))
print(f"one product, random x: relative difference of the two forms {d:.2e}")


def pcg(mul, rtol=1e-10, max_iter=20000):
    dinv = 1.0 / A.diagonal()
    xk = np.zeros_like(b); r = b.copy(); z = dinv * r; p = z.copy(); g = r @ z; nb = np.linalg.norm(b)
    for k in range(1, max_iter + 1):
        w = mul(p)
        alpha = g / (p @ w)
        xk += alpha * p; r -= alpha * w
        if np.linalg.norm(r) <= rtol * nb:
            return xk, k
        z = dinv * r; g2 = r @ z; p = z + (g2 / g) * p; g = g2
    return xk, max_iter


for name, mul in (("stored diagonal ", mul_stored), ("difference form ", mul_difference)):
    xk, k = pcg(mul)
    print(f"{name}: {k} iterations, true residual {np.linalg.norm(b - A @ xk) / np.linalg.norm(b):.2e}, "
          f"vs spsolve {np.max(np.abs(xk - v_ref[free])) / np.max(np.abs(v_ref[free])):.2e}")
