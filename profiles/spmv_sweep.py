"""Time the solver's dominant kernel on a BASELINE config (run on the GPU box).

    python profiles/spmv_sweep.py [config] [grid]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from bench import make_board  # noqa: E402
from padne_b200 import capi  # noqa: E402
from padne_b200.stamps import build_stamps  # noqa: E402

config = sys.argv[1] if len(sys.argv) > 1 else "2layer_1M"
grid = int(sys.argv[2]) if len(sys.argv) > 2 else None
fs = make_board(config, grid).to_flat()
st = build_stamps(fs)
ctx = capi.Context(0)
nnz = ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
v, info = ctx.solve(st[4], n_nodes=fs.n_nodes)
indptr = ctx.get_csr()[0]
alg = 12.0 * int(indptr[fs.n_nodes]) + 20.0 * fs.n_nodes
out = {"config": config, "N": fs.N, "nnz": nnz, "iterations": info.iterations, "t_krylov_ms": info.t_krylov_ms,
       "us_per_iteration": 1e3 * info.t_krylov_ms / max(info.iterations, 1), "alg_MB": alg / 1e6}
for name, mode, reps in (("warm", 0, 200), ("cold", 1, 50), ("in_situ", 2, 200)):
    ms = ctx.bench_spmv(reps, mode)
    out[name + "_us"] = ms * 1e3
    out[name + "_GBs"] = alg / (ms * 1e-3) / 1e9
print(json.dumps(out))
