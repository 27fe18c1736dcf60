"""``pdn_residual_noalloc`` (include/padne_b200.h): r = b - A x and ||r||^2 with no device allocation.

No counterpart in the reference, so the checker is the library's
EXISTING entry point for the same arithmetic: ``pdn_set_csr`` + ``pdn_spmv`` (y = A x), then b - y.  Tolerance:
the one ``test_spmv_matches_scipy`` uses (rtol 1e-12, atol 1e-9) for r, 1e-12 relative for ||r||^2.

The file name makes pytest collect it LAST (``pytest --collect-only``).  It is
self-contained -- its own matrices, the session context only for the existing entry point -- so it passes alone in
a fresh process and after the other tests.
"""
import numpy as np
import pytest
import scipy.sparse

from padne_b200 import capi

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-12, 1e-9


def grid_system(n, seed):
    """5-point stencil on an n x n grid with random positive edge weights plus one long coupling row: a matrix of
    the hot path's shape (about 5 entries per row, one row with hundreds) without going through the assembly."""
    rng = np.random.default_rng(seed)
    idx = np.arange(n * n).reshape(n, n)
    ii = np.concatenate([idx[:, :-1].ravel(), idx[:-1, :].ravel()])
    jj = np.concatenate([idx[:, 1:].ravel(), idx[1:, :].ravel()])
    w = rng.uniform(0.5, 2.0, ii.size)
    N = n * n + 1
    A = scipy.sparse.coo_matrix((np.concatenate([w, w]), (np.concatenate([ii, jj]), np.concatenate([jj, ii]))),
                                shape=(N, N)).tocsr()
    A = A - scipy.sparse.diags(np.asarray(A.sum(axis=1)).ravel())
    long_cols = rng.choice(n * n, size=min(300, n * n), replace=False)
    extra = scipy.sparse.coo_matrix((rng.standard_normal(long_cols.size), (np.full(long_cols.size, N - 1), long_cols)),
                                    shape=(N, N))
    A = (A + extra + extra.T).tocsr()
    A.sort_indices()
    return A


def device_csr(torch, A):
    dev = torch.device("cuda", 0)
    return (torch.from_numpy(A.indptr.astype(np.int32)).to(dev), torch.from_numpy(A.indices.astype(np.int32)).to(dev),
            torch.from_numpy(A.data.astype(np.float64)).to(dev))


@pytest.mark.parametrize("n", [37, 300, 1100])      # 2 CTAs / 88 CTAs / 1182 CTAs (more partials than threads)
def test_residual_noalloc_matches_spmv_entry_point_twice(ctx, n):
    import torch
    A = grid_system(n, seed=n)
    N = A.shape[0]
    indptr, indices, data = device_csr(torch, A)
    ctx.set_csr(indptr, indices, data)                       # the existing entry point works on the same matrix
    rng = np.random.default_rng(1)
    b = torch.from_numpy(rng.standard_normal(N)).cuda()
    r = torch.empty(N, dtype=torch.float64, device="cuda")   # outputs are the caller's
    rr = torch.empty(1, dtype=torch.float64, device="cuda")
    seen = []
    for call in range(2):                                    # two calls in a row share the scratch array
        x = torch.from_numpy(rng.standard_normal(N)).cuda()
        r.fill_(float("nan"))
        rr.fill_(float("nan"))
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0] if call == 1 else None     # (the first call may load the module)
        capi.residual_noalloc(indptr, indices, data, x, b, r, rr, device=0, stream=0)
        torch.cuda.synchronize()
        if call == 1:
            assert torch.cuda.mem_get_info()[0] == free0     # no device memory was taken
        r_ref = b - ctx.spmv(x)                              # pdn_spmv: the existing entry point
        r_np, ref_np = r.cpu().numpy(), r_ref.cpu().numpy()
        assert np.isfinite(r_np).all()
        assert np.allclose(r_np, ref_np, rtol=RTOL, atol=ATOL)
        assert np.allclose(r_np, b.cpu().numpy() - A @ x.cpu().numpy(), rtol=RTOL, atol=ATOL)
        rr_ref = float(np.dot(ref_np, ref_np))
        assert abs(float(rr.item()) - rr_ref) <= 1e-12 * rr_ref
        seen.append((r_np.copy(), float(rr.item()), x))
    assert not np.array_equal(seen[0][0], seen[1][0])        # the second call did its own work
    # the same inputs again give the same bits (fixed summation order, nothing left over in the scratch)
    capi.residual_noalloc(indptr, indices, data, seen[1][2], b, r, rr)
    torch.cuda.synchronize()
    assert np.array_equal(r.cpu().numpy(), seen[1][0]) and float(rr.item()) == seen[1][1]


def test_residual_noalloc_limits_and_errors():
    import torch
    z32 = torch.zeros(2, dtype=torch.int32, device="cuda")
    z64 = torch.zeros(2, dtype=torch.float64, device="cuda")
    with pytest.raises(NotImplementedError, match="PDN_NOALLOC_MAX_ROWS"):
        capi.residual_noalloc(z32, z32, z64, z64, z64, z64, z64, n_rows=capi.NOALLOC_MAX_ROWS + 1)
    with pytest.raises(ValueError, match="negative"):
        capi.residual_noalloc(z32, z32, z64, z64, z64, z64, z64, n_rows=-1)
    with pytest.raises(ValueError, match="device"):
        capi.residual_noalloc(z32, z32, z64, z64, z64, z64, z64, device=63)
    rr = torch.full((1,), 7.0, dtype=torch.float64, device="cuda")
    capi.residual_noalloc(z32, z32, z64, z64, z64, z64, rr, n_rows=0)       # empty system: ||r||^2 = 0
    torch.cuda.synchronize()
    assert float(rr.item()) == 0.0
