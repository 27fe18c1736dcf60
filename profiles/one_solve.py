"""One assembly + one short two-level solve (max_iter = 32): the launch list of the SETUP kernels under ncu.

    ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file out.csv python profiles/one_solve.py
"""
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
    ctx.solve(st[4], n_nodes=fs.n_nodes, max_iter=int(os.environ.get("ONE_SOLVE_ITERS", "32")))
except capi.NotConvergedError as e:
    print("expected:", str(e)[:80])
