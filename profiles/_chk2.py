import sys, os, numpy as np
sys.path.insert(0, os.getcwd())
from bench import make_board
from padne_b200 import capi
from padne_b200.stamps import build_stamps
fs = make_board("2layer_1M", int(sys.argv[1])).to_flat()
st = build_stamps(fs)
ctx = capi.Context(0)
ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
for rtol in (1e-2, 1e-4, 1e-6):
    try:
        v1, i1 = ctx.solve(st[4], n_nodes=fs.n_nodes, rtol=rtol, check_every=int(sys.argv[2]) if len(sys.argv) > 2 else 0)
        v2, i2 = ctx.solve(st[4], n_nodes=fs.n_nodes, rtol=rtol, check_every=int(sys.argv[2]) if len(sys.argv) > 2 else 0)
        print(fs.N, rtol, i1.iterations, i2.iterations, i1.restarts, i1.rel_residual, i2.rel_residual, np.array_equal(v1, v2))
    except Exception as e:
        print(fs.N, rtol, "FAIL", str(e)[-120:])
