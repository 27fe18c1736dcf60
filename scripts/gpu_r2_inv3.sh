# coarse inverse: phase counters + A/B only (no test suite)
bash scripts/gpu_r2_invt.sh 2>&1 | grep -E "inv b=|t_coarse" | tail -7
for cfg in "v1 16" "v2 12" "v2 16"; do
set -- $cfg
PADNE_B200_INVERT=$1 PADNE_B200_CO_R=$2 timeout 600 python - <<'PY' 2>&1 | tail -3
import hashlib, os, sys, numpy as np
sys.path.insert(0, "tests")
from padne_b200 import capi, synth
from test_gpu_parity import gpu_assemble
ctx = capi.Context(0)
for name, fs in (("2layer_1M", synth.board_2layer(708, n_via=16).to_flat()), ("8layer g300", synth.board_8layer(300, n_via=2, size_mm=30.0, seed=4).to_flat()),
                 ("regulator g200", synth.board_4layer_regulator(200, n_via=4, size_mm=30.0, seed=5).to_flat())):
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    best = None
    for rep in range(4):
        v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
        best = info.t_coarse_ms if best is None else min(best, info.t_coarse_ms)
    print(os.environ["PADNE_B200_INVERT"], "R", os.environ["PADNE_B200_CO_R"], name, "nc", info.n_coarse, "t_coarse_ms", round(best, 3), "iters", info.iterations,
          "res", info.residual_norm, "sha", hashlib.sha1(np.asarray(v).tobytes()).hexdigest()[:12])
PY
done
