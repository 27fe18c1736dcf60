import sys, os, numpy as np, scipy.sparse
sys.path.insert(0, os.getcwd())
from bench import make_board
from padne_b200 import capi
from padne_b200.stamps import build_stamps
fs = make_board("2layer_1M", int(sys.argv[1]) if len(sys.argv)>1 else 400).to_flat()
st = build_stamps(fs)
ctx = capi.Context(0)
ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
ip, ix, da = ctx.get_csr()
L = scipy.sparse.csr_matrix((da, ix, ip), shape=(fs.N, fs.N))
x = np.random.default_rng(0).standard_normal(fs.N)
for rep in range(3):
    y = ctx.spmv(x)
    ref = L @ x
    err = np.abs(y - ref)
    bad = np.nonzero(err > 1e-6 * np.abs(ref).max())[0]
    print(os.environ.get("PADNE_B200_LIB"), fs.N, "max err", err.max(), "bad rows", len(bad), bad[:10], (bad[:10] // 128))
