"""Design check for the next multi-GPU step, numpy/scipy on the CPU.

Today's partitioned PCG synchronises three times per iteration (dots after the SpMV, the coarse residual after
the update, the halo before the SpMV).  The candidate keeps ONE reduction per iteration:

* Chronopoulos-Gear recurrences: z = M r, w = A z, then gamma = (r, z) and delta = (w, z) in one reduction;
  beta = gamma / gamma_old, alpha = gamma / (delta - beta gamma / alpha_old); p = z + beta p, s = w + beta s,
  x += alpha p, r -= alpha s;
* the coarse residual of the two-level preconditioner M = D^-1 + P Ac^-1 P^T is not re-restricted from r:
  c_w = P^T w is formed in the SpMV tail and summed in the SAME reduction as the dots, and
  c_s = c_w + beta c_s, r_c -= alpha c_s follow r's own recurrence.

Checked here on the reference's matrix (oracle assembly) of a 2-layer board with Dirichlet rows eliminated:
iterations to 1e-10 against textbook PCG with the same preconditioner, the drift of the recurred r_c against
P^T r, and the final true residual.

    python scripts/proto_single_sync_pcg.py [grid_side] [bins_per_side]
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

side = int(sys.argv[1]) if len(sys.argv) > 1 else 224
bins = int(sys.argv[2]) if len(sys.argv) > 2 else 20

board = synth.board_2layer(side, n_via=4, size_mm=100.0, seed=0)
fs = board.to_flat()
st = build_stamps(fs)
indptr, indices, data = oracle.assemble(fs, st[:4])
L = sp.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
rhs = st[4]
v_ref = spla.spsolve(L.tocsc(), rhs)

# reduced SPD system: node block of -L with the source terminals (known voltages) eliminated
n = fs.n_nodes
vs = np.flatnonzero(fs.e_kind == 2)
fixed = np.unique(np.concatenate([fs.e_node[vs, 0], fs.e_node[vs, 1], [fs.i_gnd]]))
free = np.setdiff1d(np.arange(n), fixed)
A_full = (-L[:n, :n]).tocsr()
A = A_full[free][:, free].tocsr()
b = -rhs[:n][free] - A_full[free][:, fixed] @ v_ref[fixed]
x_ref = v_ref[free]

# two-level preconditioner: Jacobi + piecewise constants over coordinate bins per layer
xy = fs.xy[free] if n == fs.n_vertices else np.vstack([fs.xy, np.zeros((n - fs.n_vertices, 2))])[free]
mesh_of = np.searchsorted(fs.v_off, np.minimum(free, fs.n_vertices - 1), side="right") - 1
span = xy.max(0) - xy.min(0) + 1e-9
cell = np.minimum((bins * (xy - xy.min(0)) / span).astype(int), bins - 1)
agg = np.unique(mesh_of * bins * bins + cell[:, 0] * bins + cell[:, 1], return_inverse=True)[1]
nc = agg.max() + 1
P = sp.csr_matrix((np.ones(len(free)), (np.arange(len(free)), agg)), shape=(len(free), nc))
Ac_inv = np.linalg.inv((P.T @ A @ P).toarray())
dinv = 1.0 / A.diagonal()


def M(r, rc=None):
    rc = P.T @ r if rc is None else rc
    return dinv * r + P @ (Ac_inv @ rc)


def pcg(rtol=1e-10, max_iter=5000):
    x = np.zeros_like(b); r = b.copy(); z = M(r); p = z.copy(); gamma = r @ z; nb = np.linalg.norm(b)
    for k in range(1, max_iter + 1):
        w = A @ p
        alpha = gamma / (p @ w)
        x += alpha * p; r -= alpha * w
        if np.linalg.norm(r) <= rtol * nb:
            return x, k
        z = M(r); g2 = r @ z; p = z + (g2 / gamma) * p; gamma = g2
    return x, max_iter


def pcg_single_sync(rtol=1e-10, max_iter=5000):
    x = np.zeros_like(b); r = b.copy(); nb = np.linalg.norm(b)
    rc = P.T @ r                      # restricted once; afterwards by recurrence
    p = np.zeros_like(b); s = np.zeros_like(b); cs = np.zeros(nc)
    gamma_old = alpha_old = 1.0
    drift = 0.0
    for k in range(1, max_iter + 1):
        z = M(r, rc)
        w = A @ z
        gamma, delta, cw, rr = r @ z, w @ z, P.T @ w, r @ r      # ONE reduction: 3 scalars + nc coarse sums
        if np.sqrt(rr) <= rtol * nb:
            return x, k - 1, drift
        beta = 0.0 if k == 1 else gamma / gamma_old
        alpha = gamma / delta if k == 1 else gamma / (delta - beta * gamma / alpha_old)
        p = z + beta * p; s = w + beta * s; cs = cw + beta * cs
        x += alpha * p; r -= alpha * s; rc -= alpha * cs
        drift = max(drift, np.linalg.norm(rc - P.T @ r) / max(np.linalg.norm(P.T @ r), 1e-300))
        gamma_old, alpha_old = gamma, alpha
    return x, max_iter, drift


x1, k1 = pcg()
x2, k2, drift = pcg_single_sync()
nb = np.linalg.norm(b)
rel = lambda x: np.max(np.abs(x - x_ref)) / np.max(np.abs(x_ref))
print(f"rows {A.shape[0]}  coarse {nc}")
print(f"textbook PCG      : {k1} iterations, true residual {np.linalg.norm(b - A @ x1) / nb:.2e}, vs spsolve {rel(x1):.2e}")
print(f"single-reduction  : {k2} iterations, true residual {np.linalg.norm(b - A @ x2) / nb:.2e}, vs spsolve {rel(x2):.2e}, "
      f"max relative drift of recurred r_c {drift:.2e}")
