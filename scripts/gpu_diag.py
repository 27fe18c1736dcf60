"""GPU-box diagnostic (not a test): SolverInfo of every golden system and a few boards through the C ABI, one JSON
line per case -- residual_norm, ground current against the reference's, iterations, the bar the loop ran to.

    python scripts/gpu_diag.py [tag]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

from conftest import golden_system_files, load_system  # noqa: E402
from padne_b200 import capi, synth  # noqa: E402
from padne_b200.stamps import build_stamps  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "default"
ctx = capi.Context(0)


def run(name, fs, v_ref=None, ig_ref=None, res_ref=None, **kw):
    st = build_stamps(fs)
    ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
    t0 = time.perf_counter()
    try:
        v, info = ctx.solve(st[4], n_nodes=fs.n_nodes, **kw)
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"tag": tag, "case": name, "N": fs.N, "error": f"{type(e).__name__}: {e}"}), flush=True)
        return
    dt = time.perf_counter() - t0
    out = {"tag": tag, "case": name, "N": fs.N, "method": info.method, "kind": info.iteration_kind, "precond": info.precond,
           "iters": info.iterations, "restarts": info.restarts, "stagnated": info.stagnated,
           "rel_residual": info.rel_residual, "residual_norm": info.residual_norm, "ground": info.ground_node_current,
           "abs_target": info.abs_target, "floor_abs": info.floor_abs, "rhs_norm": info.rhs_norm, "nc": info.n_coarse,
           "halo": info.spmv_halo, "coverage": round(info.spmv_window_coverage, 4), "t_krylov_ms": round(info.t_krylov_ms, 3),
           "t_setup_ms": round(info.t_setup_ms, 3), "t_coarse_ms": round(info.t_coarse_ms, 3), "wall_ms": round(1e3 * dt, 2)}
    if v_ref is not None:
        n = fs.n_nodes
        out["v_err"] = float(np.max(np.abs(v[:n] - v_ref[:n])) / (np.max(np.abs(v_ref[:n])) or 1.0))
        out["ground_ref"] = float(ig_ref)
        out["residual_ref"] = float(res_ref)
    print(json.dumps(out), flush=True)


for path in golden_system_files():
    fs, d = load_system(path)
    run(os.path.basename(path), fs, d["v"], d["ground_node_current"], d["residual_norm"])
boards = [("2layer_160", synth.board_2layer(160, n_via=5, size_mm=25.0, seed=21)),
          ("regulator_100", synth.board_4layer_regulator(100, n_via=4, size_mm=30.0, seed=5)),
          ("8layer_60", synth.board_8layer(60, n_via=3, size_mm=40.0, seed=2)),
          ("tht_standin_48", synth.board_via_tht_standin(48, seed=3)),
          ("2layer_300", synth.board_2layer(300, n_via=6, size_mm=40.0, seed=13)),
          ("2layer_708", synth.board_2layer(708, n_via=16)),
          ("regulator_400", synth.board_4layer_regulator(400, n_via=8))]
for name, b in boards:
    run(name, b.to_flat())
run("2layer_708_rtol_only", boards[5][1].to_flat(), atol=-1.0)
