# per-phase cycle counters of the coarse inverse (library built with -DPDN_INV_TIMING=1 as libpadne_b200_invt.so)
PADNE_B200_LIB=padne_b200/_lib/libpadne_b200_invt.so PADNE_B200_TRACE=1 timeout 600 python - <<'PY' 2>&1 | grep -E "inv b=|trace|t_coarse" | tail -24
import sys, numpy as np
sys.path.insert(0, "tests")
from padne_b200 import capi, synth
from test_gpu_parity import gpu_assemble
ctx = capi.Context(0)
fs = synth.board_2layer(708, n_via=16).to_flat()
indptr, indices, data, rhs = gpu_assemble(ctx, fs)
for rep in range(2):
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
print("t_coarse_ms", info.t_coarse_ms, "nc", info.n_coarse)
PY
