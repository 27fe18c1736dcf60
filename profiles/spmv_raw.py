"""Time the SpMV kernel on the assembled FULL system of a config without solving (kernel experiments)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import make_board
from padne_b200 import capi
from padne_b200.stamps import build_stamps
config = sys.argv[1] if len(sys.argv) > 1 else "2layer_1M"
grid = int(sys.argv[2]) if len(sys.argv) > 2 else None
fs = make_board(config, grid).to_flat()
st = build_stamps(fs)
ctx = capi.Context(0)
nnz = ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
alg = 12.0 * nnz + 20.0 * fs.N
out = {"lib": os.environ.get("PADNE_B200_LIB", "default"), "N": fs.N, "nnz": nnz}
for name, mode, reps in (("warm", 0, 100), ("cold", 1, 30), ("in_situ", 2, 100)):
    ms = ctx.bench_spmv(reps, mode)
    out[name + "_us"] = round(ms * 1e3, 1)
    out[name + "_GBs"] = round(alg / (ms * 1e-3) / 1e9)
print(json.dumps(out))
