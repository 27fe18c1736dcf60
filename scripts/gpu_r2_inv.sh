# coarse inverse A/B: round-1 kernel (PADNE_B200_INVERT=v1) against the round-2 one, same board, then the GPU suite
mkdir -p gpurun_out
for inv in v1 v2; do
PADNE_B200_INVERT=$inv timeout 600 python - <<'PY' 2>&1 | tail -4
import hashlib, os, numpy as np, torch
from padne_b200 import capi, synth
ctx = capi.Context(0)
for name, fs in (("2layer_1M", synth.board_2layer(708, n_via=16).to_flat()), ("8layer g300", synth.board_8layer(300, n_via=2, size_mm=30.0, seed=4).to_flat()),
                 ("regulator g200", synth.board_4layer_regulator(200, n_via=4, size_mm=30.0, seed=5).to_flat())):
    from padne_b200.stamps import build_rhs, build_stamps
    import sys; sys.path.insert(0, "tests")
    from test_gpu_parity import gpu_assemble
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    best = None
    for rep in range(4):
        v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
        best = info.t_coarse_ms if best is None else min(best, info.t_coarse_ms)
    print(os.environ["PADNE_B200_INVERT"], name, "nc", info.n_coarse, "t_coarse_ms", round(best, 3), "iters", info.iterations, "res", info.residual_norm,
          "sha", hashlib.sha1(np.asarray(v).tobytes()).hexdigest()[:12])
PY
done
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-named-configs 2>/dev/null | grep '^{' | python -c "
import sys,json
d=json.loads(sys.stdin.read()); c=d['config']; print(round(d['value']/1e6,2),'MDOF/s',round(d['ms_per_step'],2),'ms iters',c['iterations'],'t_coarse',round(c['t_coarse_ms'],3),'setup',round(c['t_setup_ms'],2),'us/it',round(c['us_per_iteration'],1))"
