"""The reference-facing host API (``padne_b200.solver``) on the GPU -- reads like the reference's tests."""
import os

import numpy as np
import pytest
import scipy.sparse

from conftest import golden_system_files, load_system
from padne_b200 import solver as S
from padne_b200 import synth

pytestmark = pytest.mark.gpu


class _P:
    def __init__(self, x, y):
        self.x, self.y = x, y


class _V:
    def __init__(self, i, x, y):
        self.i, self.p = i, _P(x, y)


class _E:
    def __init__(self, origin):
        self.origin = origin


class _F:
    def __init__(self, vs):
        self.edges = [_E(v) for v in vs]


class DuckMesh:
    """Anything with .vertices / .faces like padne.mesh.Mesh (mesh.py:148-160, 203-205)."""
    def __init__(self, pts, tris):
        self.vertices = [_V(i, x, y) for i, (x, y) in enumerate(pts)]
        self.faces = [_F([self.vertices[i] for i in t]) for t in tris]


def test_laplace_operator_unit_square():
    # tests/test_solver.py:776-852
    msh = DuckMesh([(0, 0), (1, 0), (1, 1), (0, 1), (0.5, 0.5)], [(0, 1, 4), (1, 2, 4), (2, 3, 4), (3, 0, 4)])
    L = S.laplace_operator(msh)
    assert isinstance(L, scipy.sparse.coo_matrix) and L.shape == (5, 5) and L.nnz == 13
    D = L.toarray()
    assert np.array_equal(D[4], [1, 1, 1, 1, -4])
    for i in range(4):
        assert D[i, 4] == 1 and D[i, i] == -1
    # assert_matrix_is_laplacian (tests/test_solver.py:757-773)
    assert (np.diag(D) < 0).all() and np.array_equal(D, D.T) and np.allclose(D.sum(axis=1), 0)


@pytest.mark.parametrize("path", golden_system_files()[:2], ids=os.path.basename)
def test_assemble_flat_and_solve_system(path):
    fs, d = load_system(path)
    L, r = S.assemble_flat(fs)
    assert L.shape == (fs.N, fs.N) and L.nnz == len(d["data"])
    assert np.array_equal(L.indptr, d["indptr"]) and np.array_equal(L.indices, d["indices"])
    assert np.array_equal(L.data, d["data"]) and np.array_equal(r, d["r"])
    assert L.tocsc().shape == L.shape                                   # benchmarks.py:652-654
    v, info = S.solve_system(L, r)                                       # resident fast path
    assert np.max(np.abs(v - d["v"])) <= 1e-6 * np.max(np.abs(d["v"]))
    assert info.residual_norm < 1e-6 and abs(info.ground_node_current - d["ground_node_current"]) < 1e-7
    # a plain scipy matrix (e.g. the reference's own lil_matrix) takes the upload path
    v2, info2 = S.solve_system(scipy.sparse.lil_matrix(L), r)
    assert np.max(np.abs(v2 - d["v"])) <= 1e-6 * np.max(np.abs(d["v"]))


def test_solve_system_singular_behaves_like_superlu():
    b = synth.board_2layer(24, n_via=0, size_mm=10.0, seed=2)
    b.add_voltage_source(5, 6, 0.0)
    b.add_voltage_source(6, 7, 0.0)
    b.add_voltage_source(7, 5, 0.0)
    L, r = S.assemble_flat(b.to_flat())
    with pytest.warns(Warning, match="exactly singular"):
        v, info = S.solve_system(L, r)
    assert np.isnan(v).all()
