"""Iterations / time of the Krylov solve per preconditioner (run on the GPU box).

    python profiles/precond_sweep.py [config] [grid] [n_coarse ...]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from bench import make_board  # noqa: E402
from padne_b200 import capi  # noqa: E402
from padne_b200.stamps import build_stamps  # noqa: E402

config = sys.argv[1] if len(sys.argv) > 1 else "2layer_1M"
grid = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2] != "full" else None
coarse = [int(a) for a in sys.argv[3:]] or [256, 512, 1024, 1536]
fs = make_board(config, grid).to_flat()
st = build_stamps(fs)
ctx = capi.Context(0)
ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
v0 = None
for name, kw in [("jacobi", dict(precond=capi.PRECOND_JACOBI))] + [(f"two-level/{c}", dict(precond=capi.PRECOND_TWOLEVEL, n_coarse=c)) for c in coarse]:
    try:
        for rep in range(2):
            v, info = ctx.solve(st[4], n_nodes=fs.n_nodes, **kw)
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"config": config, "precond": name, "error": str(e)[-300:]}))
        continue
    if v0 is None:
        v0 = v.copy()
    diff = float(np.max(np.abs(v[:fs.n_nodes] - v0[:fs.n_nodes])) / np.max(np.abs(v0[:fs.n_nodes])))
    print(json.dumps({"config": config, "N": fs.N, "precond": info.precond, "n_coarse": info.n_coarse,
                      "iterations": info.iterations, "restarts": info.restarts, "t_krylov_ms": round(info.t_krylov_ms, 2),
                      "t_setup_ms": round(info.t_setup_ms, 2), "t_coarse_ms": round(info.t_coarse_ms, 2),
                      "t_total_ms": round(info.t_total_ms, 2), "rel_residual": info.rel_residual,
                      "us_per_iteration": round(1e3 * info.t_krylov_ms / max(info.iterations, 1), 1),
                      "max_rel_diff_vs_jacobi": diff}), flush=True)
