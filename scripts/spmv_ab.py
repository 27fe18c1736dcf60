"""A/B timing of the solver's SpMV as the last solve ran it (C ABI bench hook): in situ / warm / cold, one line.

    PADNE_B200_LIB=padne_b200/_lib/libpadne_b200_<tag>.so python scripts/spmv_ab.py [config]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import make_board  # noqa: E402
from padne_b200 import capi  # noqa: E402
from padne_b200.stamps import build_stamps  # noqa: E402

config = sys.argv[1] if len(sys.argv) > 1 else "2layer_1M"
fs = make_board(config, None).to_flat()
st = build_stamps(fs)
ctx = capi.Context(0)
ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
try:
    v, info = ctx.solve(st[4], n_nodes=fs.n_nodes)
    it, us = info.iterations, 1e3 * info.t_krylov_ms / max(info.iterations, 1)
except Exception as e:  # noqa: BLE001  (debug variants may not converge)
    it, us = -1, -1.0
print(json.dumps({"lib": os.path.basename(os.environ.get("PADNE_B200_LIB", "default")), "pcg": os.environ.get("PADNE_B200_PCG", "v2"),
                  "iterations": it, "us_per_iteration": round(us, 2),
                  "spmv_in_situ_us": round(1e3 * ctx.bench_spmv(200, 2), 2), "spmv_warm_us": round(1e3 * ctx.bench_spmv(200, 0), 2),
                  "spmv_cold_us": round(1e3 * ctx.bench_spmv(50, 1), 2)}))
