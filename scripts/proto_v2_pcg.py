"""Numerics check of the round-2 PCG iteration, numpy on the CPU.

The CUDA iteration keeps r_c by recurrence with c_w = P^T w restricted inside the SpMV in 32-bit FIXED POINT
(integer atomics: order-independent, so the solve stays bit-reproducible).  The scale is a true bound of |w_i|:
8 * max_diag * max|z_i| (diagonal dominance), 2^38 steps; r_c is re-restricted from r every `resync` iterations at the
same resolution.  (Round-2 finding: 2^25 steps -- or a 2^22 re-restriction -- do NOT converge when aggregates hold
16-50 rows: an error in the recurred r_c persists until the next re-restriction, and r_c is small against r.)
This script runs that iteration (quantisation included) next to textbook PCG on the reference's matrix and
reports iterations, true residuals and the worst drift of the recurred r_c.

    python scripts/proto_v2_pcg.py [grid_side] [bins_per_side] [resync]
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
resync = int(sys.argv[3]) if len(sys.argv) > 3 else 32

board = synth.board_2layer(side, n_via=4, size_mm=100.0, seed=0)
fs = board.to_flat()
st = build_stamps(fs)
indptr, indices, data = oracle.assemble(fs, st[:4])
L = sp.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
rhs = st[4]
v_ref = spla.spsolve(L.tocsc(), rhs)
n = fs.n_nodes
vs = np.flatnonzero(fs.e_kind == 2)
fixed = np.unique(np.concatenate([fs.e_node[vs, 0], fs.e_node[vs, 1], [fs.i_gnd]]))
free = np.setdiff1d(np.arange(n), fixed)
A_full = (-L[:n, :n]).tocsr()
A = A_full[free][:, free].tocsr()
b = -rhs[:n][free] - A_full[free][:, fixed] @ v_ref[fixed]
x_ref = v_ref[free]
xy = fs.xy[free]
mesh_of = np.searchsorted(fs.v_off, np.minimum(free, fs.n_vertices - 1), side="right") - 1
span = xy.max(0) - xy.min(0) + 1e-9
cell = np.minimum((bins * (xy - xy.min(0)) / span).astype(int), bins - 1)
agg = np.unique(mesh_of * bins * bins + cell[:, 0] * bins + cell[:, 1], return_inverse=True)[1]
nc = agg.max() + 1
P = sp.csr_matrix((np.ones(len(free)), (np.arange(len(free)), agg)), shape=(len(free), nc))
Ac_inv = np.linalg.inv((P.T @ A @ P).toarray()).astype(np.float32).astype(np.float64)   # applied from fp32
dinv = 1.0 / A.diagonal()
amax = A.diagonal().max()


def restrict_q(vec, fix, lim):
    q = np.clip(np.rint(vec * fix), -lim, lim)
    return (P.T @ q) / fix if fix > 0 else np.zeros(nc)


def pcg(rtol, atol=0.0, max_iter=5000):
    x = np.zeros_like(b); r = b.copy(); nb = np.linalg.norm(b)
    M = lambda r: dinv * r + P @ (Ac_inv @ (P.T @ r))
    z = M(r); p = z.copy(); gamma = r @ z
    tol = min(rtol * nb, atol) if atol > 0 else rtol * nb
    for k in range(1, max_iter + 1):
        w = A @ p
        alpha = gamma / (p @ w)
        x += alpha * p; r -= alpha * w
        if np.linalg.norm(r) <= tol:
            return x, k
        z = M(r); g2 = r @ z; p = z + (g2 / gamma) * p; gamma = g2
    return x, max_iter


def pcg_v2(rtol, atol=0.0, max_iter=5000):
    x = np.zeros_like(b); r = b.copy(); nb = np.linalg.norm(b)
    tol = min(rtol * nb, atol) if atol > 0 else rtol * nb
    p = np.zeros_like(b); s = np.zeros_like(b); cs = np.zeros(nc)
    rc = restrict_q(r, 2.0 ** 38 / np.linalg.norm(r), 2 ** 41)
    z = dinv * r + P @ (Ac_inv @ rc)
    gamma_old = alpha_old = 1.0
    drift = 0.0
    clipped = 0
    for k in range(1, max_iter + 1):
        if k > 1 and (k - 1) % resync == 0:
            rc = restrict_q(r, 2.0 ** 38 / np.sqrt(rr), 2 ** 41)     # scale of the PREVIOUS ||r||, like the kernel
        w = A @ z
        zmax = np.abs(z).max()
        fix = 2.0 ** 38 / (8.0 * amax * zmax) if zmax > 0 else 0.0
        clipped += int((np.abs(w * fix) > 2 ** 41).sum())
        cw = restrict_q(w, fix, 2 ** 41)
        gamma, delta, rr = r @ z, w @ z, r @ r
        if np.sqrt(rr) <= tol:
            return x, k - 1, drift, clipped
        beta = 0.0 if k == 1 else gamma / gamma_old
        alpha = gamma / delta if k == 1 else gamma / (delta - beta * gamma / alpha_old)
        cs = cw + beta * cs; rc = rc - alpha * cs
        p = z + beta * p; s = w + beta * s
        x += alpha * p; r -= alpha * s
        z = dinv * r + P @ (Ac_inv @ rc)
        drift = max(drift, np.linalg.norm(rc - P.T @ r) / max(np.linalg.norm(P.T @ r), 1e-300))
        gamma_old, alpha_old = gamma, alpha
    return x, max_iter, drift, clipped


nb = np.linalg.norm(b)
rel = lambda x: np.max(np.abs(x - x_ref)) / np.max(np.abs(x_ref))
print(f"rows {A.shape[0]}  coarse {nc}  ||b|| {nb:.3e}  amax {amax:.3e}  resync every {resync}")
for rtol, atol in ((1e-10, 0.0), (1e-10, 2.5e-10)):
    x1, k1 = pcg(rtol, atol)
    x2, k2, drift, clipped = pcg_v2(rtol, atol)
    print(f"rtol {rtol:g} atol {atol:g}")
    print(f"  textbook PCG : {k1} iterations, true residual abs {np.linalg.norm(b - A @ x1):.2e}, vs spsolve {rel(x1):.2e}")
    print(f"  v2 iteration : {k2} iterations, true residual abs {np.linalg.norm(b - A @ x2):.2e}, vs spsolve {rel(x2):.2e}, "
          f"max drift of r_c {drift:.2e}, clipped entries {clipped}")
