# pdn_residual_noalloc: its test alone in a fresh process, then the whole GPU suite (it is collected last),
# a bit-equality probe against pdn_spmv, and the default bench line of the final tree
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -4; echo "alone exit ${PIPESTATUS[0]}"
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4; echo "suite exit ${PIPESTATUS[0]}"
timeout 300 python - <<'PY' 2>&1 | tail -5
import sys, numpy as np, torch
sys.path.insert(0, "tests")
from test_zz_no_alloc import grid_system, device_csr
from padne_b200 import capi
ctx = capi.Context(0)
for n in (300, 1100):
    A = grid_system(n, n); N = A.shape[0]
    ip, ix, da = device_csr(torch, A); ctx.set_csr(ip, ix, da)
    g = np.random.default_rng(0)
    x = torch.from_numpy(g.standard_normal(N)).cuda(); b = torch.from_numpy(g.standard_normal(N)).cuda()
    r = torch.empty_like(x); rr = torch.empty(1, dtype=torch.float64, device="cuda")
    capi.residual_noalloc(ip, ix, da, x, b, r, rr); torch.cuda.synchronize()
    ref = b - ctx.spmv(x)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(20): capi.residual_noalloc(ip, ix, da, x, b, r, rr)
    ev1.record(); torch.cuda.synchronize()
    print(n, N, "bit-equal to b - pdn_spmv:", bool(torch.equal(r, ref)), "max abs diff", float((r - ref).abs().max()),
          "us per call", round(ev0.elapsed_time(ev1) * 50, 1))
PY
timeout 900 python bench.py > gpurun_out/r2_bench_final.log 2>&1; echo "bench exit $?"; grep '^{' gpurun_out/r2_bench_final.log | cut -c1-400
