for R in 16 12 10 8 6; do
PADNE_B200_CO_R=$R timeout 300 python - <<'PY'
import os,sys
sys.path.insert(0,'.')
from bench import make_board
from padne_b200 import capi
from padne_b200.stamps import build_stamps
fs = make_board("2layer_1M", None).to_flat(); st = build_stamps(fs)
ctx = capi.Context(0)
ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
for k in range(3):
    v, info = ctx.solve(st[4], n_nodes=fs.n_nodes)
print("R", os.environ["PADNE_B200_CO_R"], "nc", info.n_coarse, "t_coarse_ms", round(info.t_coarse_ms,3), "iters", info.iterations, "t_setup", round(info.t_setup_ms,3))
PY
done
