"""The oracle (``oracle/``) pinned against the reference's golden vectors -- CPU only.

(a) known answers from the reference's own tests: the unit-square Laplacian
    (``tests/test_solver.py:776-852``), the lumped-element stamps (``:65-147``), the power density
    of V = x (``:947-971``);
(b) fixtures under ``tests/golden/`` produced by the UNMODIFIED reference
    (``tests/golden/make_golden.py``): ``laplace_operator`` COO, ``assemble_system`` CSR, ``r``,
    ``spsolve`` voltages, per-face power density.  Everything integer is compared bit-exactly and,
    because the oracle keeps the reference's operation order, so are the fp64 values.
"""
import os

import numpy as np
import pytest
import scipy.sparse

from conftest import golden_laplace_files, golden_system_files, load_system
from oracle import oracle
from padne_b200.flat import FlatSystem, KIND_CURRENT, KIND_REGULATOR, KIND_RESISTOR, KIND_VOLTAGE
from padne_b200.stamps import build_stamps


def _single_mesh_system(xy, tri, conductance=1.0):
    xy = np.asarray(xy, np.float64).reshape(-1, 2)
    tri = np.asarray(tri, np.int32).reshape(-1, 3)
    return FlatSystem(v_off=[0, len(xy)], t_off=[0, len(tri)], xy=xy, tri=tri, conductance=[conductance])


def oracle_laplace(xy, tri):
    """laplace_operator through the oracle: one mesh, conductance 1, no stamps, no ground row."""
    fs = _single_mesh_system(xy, tri)
    n = fs.n_vertices
    z = np.zeros(0, np.int32)
    lib = oracle._load()
    import ctypes
    h, nnz = ctypes.c_void_p(), ctypes.c_int64()
    rc = lib.oracle_assemble(1, oracle._p(fs.v_off), oracle._p(fs.xy), oracle._p(fs.t_off), oracle._p(fs.tri),
                             oracle._p(fs.conductance), n, 0, oracle._p(z), oracle._p(z), oracle._p(np.zeros(0)),
                             oracle._p(z), ctypes.byref(h), ctypes.byref(nnz))
    if rc:
        raise oracle.OracleError(rc)
    indptr, indices, data = np.empty(n + 1, np.int32), np.empty(nnz.value, np.int32), np.empty(nnz.value)
    lib.oracle_system_csr(h, oracle._p(indptr), oracle._p(indices), oracle._p(data))
    lib.oracle_system_free(h)
    return indptr, indices, data


def test_unit_square_known_answer():
    # tests/test_solver.py:776-852: centre row [1,1,1,1,-4]; corner rows L[i,4]=1, L[i,i]=-1; cot 90 = 0
    xy = [(0, 0), (1, 0), (1, 1), (0, 1), (0.5, 0.5)]
    tri = [(0, 1, 4), (1, 2, 4), (2, 3, 4), (3, 0, 4)]
    indptr, indices, data = oracle_laplace(xy, tri)
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(5, 5)).toarray()
    expect = np.zeros((5, 5))
    expect[4, :4] = 1
    expect[:4, 4] = 1
    expect[4, 4] = -4
    for i in range(4):
        expect[i, i] = -1
    assert np.array_equal(L, expect)
    assert len(data) == 13          # corner-corner entries are absent, not stored zeros


@pytest.mark.parametrize("path", golden_laplace_files(), ids=os.path.basename)
def test_laplace_golden(path):
    d = np.load(path)
    indptr, indices, data = oracle_laplace(d["xy"], d["tri"])
    assert np.array_equal(indptr, d["indptr"])
    assert np.array_equal(indices, d["indices"])
    assert np.array_equal(data, d["data"])            # bit-exact: same operation order, no FMA
    # the reference's COO (emission order) describes the same matrix
    n = len(d["xy"])
    coo = scipy.sparse.coo_matrix((d["coo_data"], (d["coo_row"], d["coo_col"])), shape=(n, n)).tocsr()
    coo.sort_indices()
    assert np.array_equal(coo.indices, indices) and np.array_equal(coo.data, data)


def test_laplace_properties():
    # assert_matrix_is_laplacian, tests/test_solver.py:757-773
    d = np.load(golden_laplace_files()[1])
    indptr, indices, data = oracle_laplace(d["xy"], d["tri"])
    n = len(d["xy"])
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(n, n)).toarray()
    assert (np.diag(L) < 0).all()
    off = L - np.diag(np.diag(L))
    assert (off >= 0).all()
    assert np.allclose(L, L.T, atol=0)
    assert np.allclose(L.sum(axis=1), 0, atol=1e-9)


@pytest.mark.parametrize("path", golden_system_files(), ids=os.path.basename)
def test_system_golden(path):
    fs, d = load_system(path)
    stamps = build_stamps(fs)
    indptr, indices, data = oracle.assemble(fs, stamps)
    assert np.array_equal(indptr, d["indptr"])
    assert np.array_equal(indices, d["indices"])
    assert np.array_equal(data, d["data"])
    assert np.array_equal(stamps[4], d["r"])
    v, res, ig = oracle.solve(indptr, indices, data, stamps[4])
    assert np.allclose(v, d["v"], rtol=1e-9, atol=1e-12)
    assert res < 1e-8


@pytest.mark.parametrize("path", golden_system_files()[:3], ids=os.path.basename)
def test_power_density_golden(path):
    fs, d = load_system(path)
    cond = d["layer_conductance"][d["mesh_layer"]]
    got = []
    for m in range(fs.n_mesh):
        v0, v1 = fs.v_off[m], fs.v_off[m + 1]
        t0, t1 = fs.t_off[m], fs.t_off[m + 1]
        got.append(oracle.power_density(fs.xy[v0:v1], d["v"][v0:v1], fs.tri[t0:t1], cond[m]))
    assert np.array_equal(np.concatenate(got), d["power_density"])


def test_power_density_known_answer():
    # tests/test_solver.py:947-971: V = x, sigma = 2 -> p = 2.0 on every face
    xy = np.array([(0, 0), (1, 0), (1, 1), (0, 1)], np.float64)
    tri = np.array([(0, 1, 2), (0, 2, 3)], np.int32)
    p = oracle.power_density(xy, xy[:, 0].copy(), tri, 2.0)
    assert np.allclose(p, 2.0, atol=1e-12)
    p = oracle.power_density(xy, np.full(4, 3.3), tri, 2.0)
    assert np.array_equal(p, np.zeros(2))


def _lumped(kinds, nodes, vals, gains, n_internal, n_extra, xidx):
    z = np.zeros(1, np.int64)
    return FlatSystem(v_off=z, t_off=z, xy=np.zeros((0, 2)), tri=np.zeros((0, 3), np.int32),
                      conductance=np.zeros(0), n_internal=n_internal, n_extra=n_extra, i_gnd=0,
                      e_kind=kinds, e_node=nodes, e_val=vals, e_gain=gains, e_xidx=xidx)


def _dense_solve_drop_ground(fs):
    # TestNetworkSolver (tests/test_solver.py:56-58): dense solve with row/column 0 dropped as ground
    s = build_stamps(fs, with_ground=False)
    N = fs.N - 1                                    # no ground row in those tests
    indptr, indices, data = oracle.assemble(_resize(fs), s)
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N)).toarray()[:N, :N]
    r = s[4][:N]
    v = np.zeros(N)
    v[1:] = np.linalg.solve(L[1:, 1:], r[1:])
    return v


def _resize(fs):
    return fs


def test_network_current_into_resistor():
    # tests/test_solver.py:65-83: 1.1 A through 2.2 ohm -> 2.42 V
    fs = _lumped([KIND_CURRENT, KIND_RESISTOR], [[0, 1, 0, 0], [0, 1, 0, 0]], [1.1, 2.2], [0, 0], 2, 0, [-1, -1])
    v = _dense_solve_drop_ground(fs)
    assert abs(abs(v[1] - v[0]) - 2.42) < 1e-6


def test_network_voltage_source_and_resistor():
    # tests/test_solver.py:85-107: 3.3 V across 2.2 ohm -> i = 1.5 A
    fs = _lumped([KIND_VOLTAGE, KIND_RESISTOR], [[1, 0, 0, 0], [0, 1, 0, 0]], [3.3, 2.2], [0, 0], 2, 1, [2, -1])
    v = _dense_solve_drop_ground(fs)
    assert abs((v[1] - v[0]) - 3.3) < 1e-6
    assert abs(abs(v[2]) - 1.5) < 1e-6


def test_network_regulator():
    # tests/test_solver.py:109-147: V_out = 3.3, I = V/R = 1.5, mirrored gain*I through the input resistor
    gain, r_in = 0.42, 1.0
    # nodes: 0 gnd, 1 out, 2 in ; regulator v_p=1 v_n=0 s_f=2 s_t=0 ; load 2.2 ohm on out ; r_in from in to gnd
    fs = _lumped([KIND_REGULATOR, KIND_RESISTOR, KIND_RESISTOR],
                 [[1, 0, 2, 0], [1, 0, 0, 0], [2, 0, 0, 0]], [3.3, 2.2, r_in], [gain, 0, 0], 3, 1, [3, -1, -1])
    v = _dense_solve_drop_ground(fs)
    assert abs(v[1] - 3.3) < 1e-6
    assert abs(abs(v[3]) - 1.5) < 1e-6
    assert abs(abs(v[2]) - gain * 1.5 * r_in) < 1e-6     # 0.63 V


def test_lil_zero_deletion_ldo_shape():
    # n = s_f and gain = 1.0 -> the -1 and +gain stamps cancel, entry leaves the pattern
    fs = _lumped([KIND_REGULATOR, KIND_RESISTOR], [[1, 0, 0, 2], [1, 2, 0, 0]], [3.3, 1.0], [1.0, 0], 3, 1, [3, -1])
    s = build_stamps(fs, with_ground=False)
    indptr, indices, data = oracle.assemble(fs, s)
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
    assert L[0, 3] == 0 and 3 not in indices[indptr[0]:indptr[1]]
    assert not (data == 0).any()


def test_degenerate_triangle_is_zero_division():
    with pytest.raises(oracle.OracleError) as e:
        oracle_laplace([(0, 0), (1, 0), (2, 0)], [(0, 1, 2)])
    assert e.value.code == 2


def test_corner_weights_match_laplace_offdiagonals():
    d = np.load(golden_laplace_files()[1])
    w = oracle.corner_weights(d["xy"], d["tri"])
    n = len(d["xy"])
    acc = {}
    for t, (a, b, c) in enumerate(d["tri"]):
        for corner, (i, k) in enumerate([(b, c), (c, a), (a, b)]):
            key = (min(i, k), max(i, k))
            acc.setdefault(key, []).append(w[t, corner])
    L = scipy.sparse.csr_matrix((d["data"], d["indices"], d["indptr"]), shape=(n, n))
    for (i, k), ws in acc.items():
        assert L[i, k] == sum(ws)


# ---- row F3: nearest mesh vertex (NodeIndexer's KD-tree lookup) --------------------------------------
def golden_nearest_files():
    import glob
    return sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "nearest_*.npz")))


@pytest.mark.parametrize("path", golden_nearest_files(), ids=os.path.basename)
def test_nearest_vertex_golden(path):
    # expect = what the reference's own KD-trees returned (tests/golden/make_golden_nearest.py)
    d = np.load(path)
    idx, d2 = oracle.nearest_vertex(d["v_off"], d["mesh_layer"], d["xy"], d["q_xy"], d["q_layer"])
    assert np.array_equal(idx, d["expect"])
    assert np.all(d2 >= 0.0)


def test_nearest_vertex_matches_scipy_kdtree():
    # the reference's call (padne/solver.py:390-393, 424) on random clouds; two meshes share layer 1
    import scipy.spatial
    rng = np.random.default_rng(5)
    sizes = [700, 450, 300, 1]
    mesh_layer = np.array([0, 1, 1, 3], np.int32)
    v_off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    xy = rng.uniform(-5, 5, (v_off[-1], 2))
    q_xy = rng.uniform(-6, 6, (500, 2))
    q_layer = rng.choice([0, 1, 3], 500).astype(np.int32)
    idx, d2 = oracle.nearest_vertex(v_off, mesh_layer, xy, q_xy, q_layer)
    v_layer = np.repeat(mesh_layer, sizes)
    for layer in (0, 1, 3):
        members = np.flatnonzero(v_layer == layer)
        tree = scipy.spatial.KDTree(xy[members], leafsize=32)
        sel = np.flatnonzero(q_layer == layer)
        dist, i = tree.query(q_xy[sel], k=1)
        assert np.array_equal(idx[sel], members[i])
        assert np.allclose(np.sqrt(d2[sel]), dist, rtol=1e-14, atol=0)


def test_nearest_vertex_ties_and_empty_layer():
    v_off = np.array([0, 4, 4, 6], np.int64)             # mesh 1 is empty
    mesh_layer = np.array([0, 2, 0], np.int32)
    xy = np.array([[0, 0], [1, 0], [1, 1], [0, 1], [2, 0.5], [0.5, 0.5]], float)
    q_xy = np.array([[0.5, 0.0], [0.5, 0.5], [9, 9], [0.5, 0.5]], float)
    q_layer = np.array([0, 0, 0, 2], np.int32)
    idx, d2 = oracle.nearest_vertex(v_off, mesh_layer, xy, q_xy, q_layer)
    assert idx.tolist() == [0, 5, 4, -1]                 # tie 0/1 -> lowest index; layer 2 has no vertex
    assert d2[1] == 0.0 and np.isinf(d2[3])
