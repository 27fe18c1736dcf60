# wall-clock phases of a solve incl. the CUDA-graph build (PADNE_B200_TRACE=1)
PADNE_B200_TRACE=1 timeout 600 python - <<'PY' 2>&1 | grep -E "trace|iters" | tail -12
import sys, numpy as np
sys.path.insert(0, "tests")
from padne_b200 import capi, synth
from test_gpu_parity import gpu_assemble
ctx = capi.Context(0)
fs = synth.board_2layer(708, n_via=16).to_flat()
indptr, indices, data, rhs = gpu_assemble(ctx, fs)
for rep in range(3):
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
print("iters", info.iterations, "t_krylov", info.t_krylov_ms, "t_setup", info.t_setup_ms, "t_total", info.t_total_ms)
PY
