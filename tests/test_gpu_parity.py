"""Parity of the CUDA path against the oracle and the reference's golden fixtures -- B200 only.

Everything goes through the C ABI (``padne_b200.capi``).  Bars (BASELINE.json north_star):
sparsity pattern bit-exact; fp64 matrix values bit-exact (same operation order as the reference,
no FMA); node voltages within 1e-6 relative of the reference's ``spsolve``; relative residual of
the Krylov solve <= 1e-10.
"""
import os

import numpy as np
import pytest
import scipy.sparse
import scipy.sparse.linalg

from conftest import golden_laplace_files, golden_system_files, load_system
from oracle import oracle
from padne_b200 import capi, synth
from padne_b200.flat import FlatSystem
from padne_b200.stamps import build_stamps

pytestmark = pytest.mark.gpu

VOLTAGE_RTOL = 1e-6          # north_star: voltages within 1e-6 relative of the reference
RESIDUAL_RTOL = 1e-10        # north_star: relative residual <= 1e-10


def gpu_assemble(ctx, fs):
    s_row, s_col, s_val, s_op, rhs = build_stamps(fs)
    nnz = ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, s_row, s_col, s_val, s_op)
    indptr, indices, data = ctx.get_csr()
    assert len(indices) == nnz
    return indptr, indices, data, rhs


def gpu_laplace(ctx, xy, tri):
    xy = np.ascontiguousarray(xy, np.float64)
    tri = np.ascontiguousarray(tri, np.int32)
    z = np.zeros(0, np.int32)
    ctx.assemble(np.array([0, len(xy)]), xy, np.array([0, len(tri)]), tri, np.array([1.0]), len(xy), z, z,
                 np.zeros(0), z)
    return ctx.get_csr()


def voltage_error(v, v_ref, n_nodes):
    scale = np.max(np.abs(v_ref[:n_nodes])) or 1.0
    return np.max(np.abs(v[:n_nodes] - v_ref[:n_nodes])) / scale


# ---- kernel (1) --------------------------------------------------------------------------------
@pytest.mark.parametrize("path", golden_laplace_files(), ids=os.path.basename)
def test_element_weights_bit_exact(ctx, path):
    d = np.load(path)
    w = ctx.element_weights(d["xy"], d["tri"])
    assert np.array_equal(w, oracle.corner_weights(d["xy"], d["tri"]))


def test_element_weights_large_random(ctx):
    xy, tri, _ = synth.grid_layer(300, 280, 0.13, seed=5, jitter=0.3)
    w = ctx.element_weights(xy, tri)
    assert np.array_equal(w, oracle.corner_weights(xy, tri))


def test_element_weights_empty_and_degenerate(ctx):
    assert ctx.element_weights(np.zeros((3, 2)), np.zeros((0, 3), np.int32)).shape == (0, 3)
    with pytest.raises(ZeroDivisionError):
        ctx.element_weights(np.array([(0., 0.), (1., 0.), (2., 0.)]), np.array([(0, 1, 2)], np.int32))
    with pytest.raises(ValueError):
        ctx.element_weights(np.zeros((3, 2)), np.array([(0, 1, 3)], np.int32))


# ---- kernels (1)+(2) ---------------------------------------------------------------------------
@pytest.mark.parametrize("path", golden_laplace_files(), ids=os.path.basename)
def test_laplace_operator_golden(ctx, path):
    d = np.load(path)
    indptr, indices, data = gpu_laplace(ctx, d["xy"], d["tri"])
    assert np.array_equal(indptr, d["indptr"])
    assert np.array_equal(indices, d["indices"])
    assert np.array_equal(data, d["data"])


def test_unit_square_known_answer(ctx):
    # tests/test_solver.py:776-852
    indptr, indices, data = gpu_laplace(ctx, [(0, 0), (1, 0), (1, 1), (0, 1), (0.5, 0.5)],
                                        [(0, 1, 4), (1, 2, 4), (2, 3, 4), (3, 0, 4)])
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(5, 5)).toarray()
    expect = np.zeros((5, 5))
    expect[4, :4] = expect[:4, 4] = 1
    expect[4, 4] = -4
    expect[np.arange(4), np.arange(4)] = -1
    assert np.array_equal(L, expect) and len(data) == 13


@pytest.mark.parametrize("path", golden_system_files(), ids=os.path.basename)
def test_assemble_golden(ctx, path):
    fs, d = load_system(path)
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    assert np.array_equal(indptr, d["indptr"])
    assert np.array_equal(indices, d["indices"])
    assert np.array_equal(data, d["data"])
    assert np.array_equal(rhs, d["r"])


@pytest.mark.parametrize("name,board", [
    ("2layer_200", lambda: synth.board_2layer(200, n_via=6, size_mm=30.0, seed=11)),
    ("right_angle_150", lambda: synth.board_2layer(150, n_via=4, size_mm=20.0, seed=0, jitter=0.0)),
    ("regulator_120", lambda: synth.board_4layer_regulator(120, n_via=5, size_mm=40.0, seed=3)),
    ("8layer_70", lambda: synth.board_8layer(70, n_via=3, size_mm=50.0, seed=9)),
    ("tht_standin_64", lambda: synth.board_via_tht_standin(64, seed=2)),
])
def test_assemble_matches_oracle(ctx, name, board):
    fs = board().to_flat()
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    o_indptr, o_indices, o_data = oracle.assemble(fs, build_stamps(fs))
    assert np.array_equal(indptr, o_indptr)
    assert np.array_equal(indices, o_indices)
    assert np.array_equal(data, o_data)


def test_assemble_is_deterministic(ctx):
    fs = synth.board_4layer_regulator(90, n_via=4, size_mm=30.0, seed=1).to_flat()
    a = gpu_assemble(ctx, fs)
    for _ in range(3):
        b = gpu_assemble(ctx, fs)
        assert all(np.array_equal(x, y) for x, y in zip(a[:3], b[:3]))


def test_assemble_shuffled_triangles(ctx):
    # triangle order changes Vertex.out and therefore the orbit start (summation order)
    xy, tri, _ = synth.grid_layer(60, 50, 0.2, seed=4, jitter=0.35)
    tri = tri[np.random.default_rng(0).permutation(len(tri))]
    fs = FlatSystem(v_off=[0, len(xy)], t_off=[0, len(tri)], xy=xy, tri=tri, conductance=[2082.5])
    indptr, indices, data, _ = gpu_assemble(ctx, fs)
    o = oracle.assemble(fs, build_stamps(fs))
    assert np.array_equal(indptr, o[0]) and np.array_equal(indices, o[1]) and np.array_equal(data, o[2])


def test_assemble_errors(ctx):
    z = np.zeros(0, np.int32)
    # zero-area triangle -> ZeroDivisionError (mesh.py:138)
    with pytest.raises(ZeroDivisionError):
        gpu_laplace(ctx, [(0, 0), (1, 0), (2, 0), (0, 1)], [(0, 1, 2), (0, 1, 3)])
    # two fans meeting in one vertex -> "Non-manifold mesh" (mesh.py:343)
    with pytest.raises(ValueError, match="Non-manifold"):
        gpu_laplace(ctx, [(0, 0), (1, 0), (0, 1), (-1, 0), (0, -1)], [(0, 1, 2), (0, 3, 4)])
    # triangle index out of range
    with pytest.raises(ValueError):
        gpu_laplace(ctx, [(0, 0), (1, 0), (0, 1)], [(0, 1, 5)])
    # stamp outside the matrix
    with pytest.raises(ValueError):
        ctx.assemble(np.array([0, 3]), np.array([(0., 0.), (1., 0.), (0., 1.)]), np.array([0, 1]),
                     np.array([(0, 1, 2)], np.int32), np.array([1.0]), 4, np.array([9], np.int32),
                     np.array([0], np.int32), np.array([1.0]), np.array([0], np.int32))


def test_lumped_only_system(ctx):
    # no mesh at all: TestNetworkSolver-style system (tests/test_solver.py:85-107) plus the ground row
    z = np.zeros(1, np.int64)
    fs = FlatSystem(v_off=z, t_off=z, xy=np.zeros((0, 2)), tri=np.zeros((0, 3), np.int32), conductance=np.zeros(0),
                    n_internal=2, n_extra=1, i_gnd=0, e_kind=[2, 0], e_node=[[1, 0, 0, 0], [0, 1, 0, 0]],
                    e_val=[3.3, 2.2], e_gain=[0, 0], e_xidx=[2, -1])
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    o = oracle.assemble(fs, build_stamps(fs))
    assert np.array_equal(indptr, o[0]) and np.array_equal(indices, o[1]) and np.array_equal(data, o[2])
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
    assert abs(v[1] - v[0] - 3.3) < 1e-9 and abs(abs(v[2]) - 1.5) < 1e-9
    assert abs(v[0]) < 1e-12 and abs(info.ground_node_current) < 1e-9


# ---- kernels (3)+(4) ---------------------------------------------------------------------------
@pytest.mark.parametrize("path", golden_system_files(), ids=os.path.basename)
def test_solve_golden(ctx, path):
    fs, d = load_system(path)
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)                    # defaults: rtol 1e-10 AND the reference's 1e-9 absolute bar
    assert info.method == ("bicgstab" if fs.has_regulator else "pcg")
    assert info.rel_residual <= RESIDUAL_RTOL
    assert voltage_error(v, d["v"], fs.n_nodes) <= VOLTAGE_RTOL
    # extra unknowns: source currents and the ground current (solver.py:777), recovered from node residuals
    i_scale = max(1.0, np.max(np.abs(d["v"][fs.n_nodes:])))
    assert np.allclose(v[fs.n_nodes:], d["v"][fs.n_nodes:], rtol=1e-6, atol=1e-12 * i_scale + 1e-9)
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
    assert abs(info.residual_norm - np.linalg.norm(L @ v - rhs)) <= 1e-6 * max(info.residual_norm, 1e-12) + 1e-12
    # SolverInfo parity: the reference's own bar ||L v - r||_2 < 1e-9 (tests/test_solver.py:2083-2089; its spsolve
    # leaves 1.5e-11 ... 6.2e-10 on these fixtures) and its ground current to 1e-9 A, hence no spurious SolverWarning
    assert info.residual_norm < 1e-9
    assert abs(info.ground_node_current - d["ground_node_current"]) < 1e-9
    assert np.isclose(info.ground_node_current, 0) == np.isclose(d["ground_node_current"], 0)


@pytest.mark.parametrize("name,board", [
    ("2layer_160", lambda: synth.board_2layer(160, n_via=5, size_mm=25.0, seed=21)),
    ("regulator_100", lambda: synth.board_4layer_regulator(100, n_via=4, size_mm=30.0, seed=5)),
    ("8layer_60", lambda: synth.board_8layer(60, n_via=3, size_mm=40.0, seed=2)),
    ("tht_standin_48", lambda: synth.board_via_tht_standin(48, seed=3)),
])
def test_solve_matches_spsolve(ctx, name, board):
    fs = board().to_flat()
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
    v_ref, res_ref, ig_ref = oracle.solve(indptr, indices, data, rhs)      # the reference's own spsolve call
    assert info.rel_residual <= RESIDUAL_RTOL
    assert voltage_error(v, v_ref, fs.n_nodes) <= VOLTAGE_RTOL
    # currents are recovered from node residuals; with the absolute bar they sit within 1e-8 A of spsolve's
    # (regulator_100 carries 1e4 A through 39k unknowns: its residual floor is 3e-9, see pdn_info.floor_abs)
    assert np.allclose(v[fs.n_nodes:], v_ref[fs.n_nodes:], rtol=1e-6, atol=2e-8)
    assert abs(info.ground_node_current - ig_ref) < 2e-8
    assert info.residual_norm < max(1e-9, 16 * info.floor_abs)


def test_solve_autodetects_n_nodes_and_accepts_foreign_csr(ctx):
    fs, d = load_system(golden_system_files()[0])
    ctx.set_csr(d["indptr"], d["indices"], d["data"])
    v, info = ctx.solve(d["r"])                         # n_nodes = -1 -> detected from the diagonal
    assert voltage_error(v, d["v"], fs.n_nodes) <= VOLTAGE_RTOL


def test_autodetect_with_an_internal_node_that_only_sources_touch(ctx):
    # ADVICE (round 1): an internal node between two voltage sources in series has no diagonal entry; as the LAST
    # internal node it used to be taken for a constraint row.  Small resistor ladder 0..4, node 5 internal:
    # source A: node 0 -> 5 (1 V), source B: node 5 -> 4 (0.5 V), ground at node 4.  Stamps as padne/solver.py:469-560.
    n, m = 6, 2
    N = n + m + 1
    L = scipy.sparse.lil_matrix((N, N))
    r = np.zeros(N)
    for a in range(4):                                   # conductances on the chain 0-1-2-3-4
        g = 1.0 + a
        L[a, a] -= g; L[a + 1, a + 1] -= g; L[a, a + 1] += g; L[a + 1, a] += g
    for k, (p_, q_, volt) in enumerate([(0, 5, 1.0), (5, 4, 0.5)]):
        c = n + k
        L[p_, c] += 1.0; L[q_, c] -= 1.0; L[c, p_] = 1.0; L[c, q_] = -1.0; r[c] = volt
    L[4, N - 1] = 1.0; L[N - 1, 4] = 1.0
    A = L.tocsr(); A.sort_indices()
    assert A[5, 5] == 0.0                                # the internal node stores no diagonal
    ctx.set_csr(A.indptr.astype(np.int32), A.indices.astype(np.int32), A.data)
    v, info = ctx.solve(r)                               # n_nodes = -1 -> detected
    v_ref = scipy.sparse.linalg.spsolve(A.tocsc(), r)
    assert np.allclose(v, v_ref, rtol=1e-9, atol=1e-12)
    v2, _ = ctx.solve(r, n_nodes=n)
    assert np.array_equal(v, v2)


def test_solve_voltage_source_loop_is_singular(ctx):
    # a closed loop of 0 V sources: SuperLU reports "Matrix is exactly singular" 
    b = synth.board_2layer(24, n_via=0, size_mm=10.0, seed=2)
    g0 = 5
    b.add_voltage_source(g0, g0 + 1, 0.0)
    b.add_voltage_source(g0 + 1, g0 + 2, 0.0)
    b.add_voltage_source(g0 + 2, g0, 0.0)
    fs = b.to_flat()
    _, _, _, rhs = gpu_assemble(ctx, fs)
    with pytest.raises(capi.SingularSystemError):
        ctx.solve(rhs, n_nodes=fs.n_nodes)


def test_solve_not_converged_is_loud(ctx):
    fs = synth.board_2layer(80, n_via=2, size_mm=20.0, seed=1).to_flat()
    _, _, _, rhs = gpu_assemble(ctx, fs)
    with pytest.raises(capi.NotConvergedError):
        ctx.solve(rhs, n_nodes=fs.n_nodes, max_iter=5)


def test_solve_multi_item_pipeline_matches_spsolve(ctx):
    # > 592 CTAs x 128 rows: every CTA streams several work items through the TMA pipeline
    fs = synth.board_2layer(300, n_via=6, size_mm=40.0, seed=13).to_flat()
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
    v_ref, _, _ = oracle.solve(indptr, indices, data, rhs)
    assert info.rel_residual <= RESIDUAL_RTOL
    assert voltage_error(v, v_ref, fs.n_nodes) <= VOLTAGE_RTOL
    v2, info2 = ctx.solve(rhs, n_nodes=fs.n_nodes)
    assert np.array_equal(v, v2) and info.iterations == info2.iterations


def test_solve_is_deterministic(ctx):
    fs = synth.board_2layer(100, n_via=3, size_mm=20.0, seed=7).to_flat()
    _, _, _, rhs = gpu_assemble(ctx, fs)
    v1, i1 = ctx.solve(rhs, n_nodes=fs.n_nodes)
    v2, i2 = ctx.solve(rhs, n_nodes=fs.n_nodes)
    assert np.array_equal(v1, v2) and i1.iterations == i2.iterations


# ---- kernel (4'): two-level preconditioner -----------------------------------------------------
@pytest.mark.parametrize("name,board,n_coarse", [
    ("2layer_160", lambda: synth.board_2layer(160, n_via=5, size_mm=25.0, seed=21), 0),
    ("2layer_160_c64", lambda: synth.board_2layer(160, n_via=5, size_mm=25.0, seed=21), 64),
    ("8layer_60", lambda: synth.board_8layer(60, n_via=3, size_mm=40.0, seed=2), 300),
    ("tht_standin_48", lambda: synth.board_via_tht_standin(48, seed=3), 0),       # internal nodes get aggregates
])
def test_two_level_preconditioner(ctx, name, board, n_coarse):
    from padne_b200 import capi
    fs = board().to_flat()
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    vj, ij = ctx.solve(rhs, n_nodes=fs.n_nodes, precond=capi.PRECOND_JACOBI)
    vt, it = ctx.solve(rhs, n_nodes=fs.n_nodes, precond=capi.PRECOND_TWOLEVEL, n_coarse=n_coarse)
    assert ij.precond == "jacobi" and ij.n_coarse == 0
    assert it.precond == "two-level" and it.n_coarse >= 4
    assert it.iterations < ij.iterations                     # the coarse space pays
    assert it.rel_residual <= RESIDUAL_RTOL
    v_ref, _, _ = oracle.solve(indptr, indices, data, rhs)
    assert voltage_error(vt, v_ref, fs.n_nodes) <= VOLTAGE_RTOL
    assert voltage_error(vt, vj, fs.n_nodes) <= 1e-8
    vt2, it2 = ctx.solve(rhs, n_nodes=fs.n_nodes, precond=capi.PRECOND_TWOLEVEL, n_coarse=n_coarse)
    assert np.array_equal(vt, vt2) and it.iterations == it2.iterations   # fixed-point sums: bit-reproducible


def test_two_level_needs_geometry(ctx):
    from padne_b200 import capi
    fs, d = load_system(golden_system_files()[0])
    ctx.set_csr(d["indptr"], d["indices"], d["data"])             # a foreign matrix carries no coordinates
    with pytest.raises(NotImplementedError):
        ctx.solve(d["r"], n_nodes=fs.n_nodes, precond=capi.PRECOND_TWOLEVEL)
    v, info = ctx.solve(d["r"], n_nodes=fs.n_nodes)               # AUTO falls back to Jacobi
    assert info.precond == "jacobi"


@pytest.mark.parametrize("grid,n_coarse", [(40, 0), (120, 200)])
def test_two_level_bicgstab_regulator(ctx, grid, n_coarse):
    # regulator gains make the reduced operator non-symmetric: BiCGSTAB, preconditioned twice per
    # iteration through the same (non-symmetrised) coarse inverse
    from padne_b200 import capi
    fs = synth.board_4layer_regulator(grid, n_via=2, size_mm=12.0, seed=4).to_flat()
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    vj, ij = ctx.solve(rhs, n_nodes=fs.n_nodes, precond=capi.PRECOND_JACOBI)
    vt, it = ctx.solve(rhs, n_nodes=fs.n_nodes, precond=capi.PRECOND_TWOLEVEL, n_coarse=n_coarse)
    assert ij.method == "bicgstab" and it.method == "bicgstab"
    assert it.precond == "two-level" and it.n_coarse >= 4
    assert it.iterations < ij.iterations
    assert it.rel_residual <= RESIDUAL_RTOL
    v_ref, _, _ = oracle.solve(indptr, indices, data, rhs)
    assert voltage_error(vt, v_ref, fs.n_nodes) <= VOLTAGE_RTOL
    vt2, it2 = ctx.solve(rhs, n_nodes=fs.n_nodes, precond=capi.PRECOND_TWOLEVEL, n_coarse=n_coarse)
    assert np.array_equal(vt, vt2) and it.iterations == it2.iterations
    va, ia = ctx.solve(rhs, n_nodes=fs.n_nodes)                   # AUTO picks it for n >= 1024
    assert ia.precond == "two-level"


@pytest.mark.parametrize("grid", [50, 180])       # 180: several work items per CTA (pipelined path)
def test_spmv_matches_scipy(ctx, grid):
    fs = synth.board_8layer(grid, n_via=2, size_mm=30.0, seed=4).to_flat()
    indptr, indices, data, _ = gpu_assemble(ctx, fs)
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
    x = np.random.default_rng(0).standard_normal(fs.N)
    y = ctx.spmv(x)
    assert np.allclose(y, L @ x, rtol=1e-12, atol=1e-9)


def test_full_size_properties_1M(ctx):
    # BASELINE config 2 at full size: size-independent properties instead of an oracle run
    fs = synth.board_2layer(708, n_via=16).to_flat()
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    n = fs.n_vertices
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
    Lnn = L[:n, :n]
    assert abs(Lnn - Lnn.T).max() == 0.0                              # symmetric, bit for bit
    assert (Lnn.diagonal() < 0).all()
    rows = np.abs(np.asarray(Lnn.sum(axis=1)).ravel())                # mesh rows sum to ~0 (+ via stamps sum to 0)
    assert rows.max() < 1e-6
    assert np.all(np.diff(indptr) > 0) and not (data == 0).any()
    assert all(np.all(np.diff(indices[indptr[i]:indptr[i + 1]]) > 0) for i in range(0, fs.N, 9973))
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
    assert info.rel_residual <= RESIDUAL_RTOL
    # physics: the 1 V source is honoured, the 1 A source's current returns through the ground
    src = np.nonzero(fs.e_kind == 2)[0][0]
    p, q = fs.e_node[src, 0], fs.e_node[src, 1]
    assert abs(v[p] - v[q] - 1.0) < 1e-9 and abs(info.ground_node_current) < 1e-6
    # linearity: doubling the right-hand side doubles the solution
    v2, _ = ctx.solve(2.0 * rhs, n_nodes=fs.n_nodes)
    assert voltage_error(v2, 2.0 * v, fs.n_nodes) <= 1e-8


# ---- row F1 -------------------------------------------------------------------------------------
@pytest.mark.parametrize("path", golden_system_files(), ids=os.path.basename)
def test_power_density_golden(ctx, path):
    fs, d = load_system(path)
    cond = d["layer_conductance"][d["mesh_layer"]]
    got = []
    for m in range(fs.n_mesh):
        v0, v1, t0, t1 = fs.v_off[m], fs.v_off[m + 1], fs.t_off[m], fs.t_off[m + 1]
        got.append(ctx.power_density(fs.xy[v0:v1], d["v"][v0:v1], fs.tri[t0:t1], float(cond[m])))
    assert np.array_equal(np.concatenate(got), d["power_density"])


# ---- round 2: the three-kernel PCG iteration (pdn_cg2.cuh) -------------------------------------------------------
@pytest.mark.parametrize("name,board", [
    ("2layer_160", lambda: synth.board_2layer(160, n_via=5, size_mm=25.0, seed=21)),      # 50 rows per aggregate
    ("8layer_60", lambda: synth.board_8layer(60, n_via=3, size_mm=40.0, seed=2)),
    ("tht_standin_48", lambda: synth.board_via_tht_standin(48, seed=3)),                  # internal nodes, 8 rows per aggregate
    ("2layer_300", lambda: synth.board_2layer(300, n_via=6, size_mm=40.0, seed=13)),
])
def test_round2_iteration_matches_round1(ctx, monkeypatch, name, board):
    # the coarse residual kept by recurrence (fixed-point c_w = P^T w from the SpMV) must walk the same path as the
    # round-1 iteration that restricts r afresh every iteration: same iteration count to +-3, same voltages, and
    # bit-reproducible (integer sums)
    fs = board().to_flat()
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    monkeypatch.setenv("PADNE_B200_PCG", "v1")
    v1, i1 = ctx.solve(rhs, n_nodes=fs.n_nodes)
    monkeypatch.setenv("PADNE_B200_PCG", "v2")
    v2, i2 = ctx.solve(rhs, n_nodes=fs.n_nodes)
    v2b, i2b = ctx.solve(rhs, n_nodes=fs.n_nodes)
    assert (i1.iteration_kind, i2.iteration_kind) == (1, 2) and i2.method == "pcg" and i2.precond == "two-level"
    assert abs(i2.iterations - i1.iterations) <= 3
    assert i2.rel_residual <= RESIDUAL_RTOL and i2.residual_norm < max(1e-9, 16 * i2.floor_abs)
    assert voltage_error(v2, v1, fs.n_nodes) <= 1e-9
    assert np.array_equal(v2, v2b) and i2.iterations == i2b.iterations
    v_ref, _, _ = oracle.solve(indptr, indices, data, rhs)
    assert voltage_error(v2, v_ref, fs.n_nodes) <= VOLTAGE_RTOL


@pytest.mark.parametrize("spmv", ["ring", "xs"])
def test_spmv_kernels_agree(ctx, monkeypatch, spmv):
    # the x-band staging SpMV (shared-memory window + global fallback for out-of-window columns) against the TMA-ring
    # kernel: PCG and BiCGSTAB solves land on the same voltages; a Delaunay-ordered cloud has no band at all
    from analytic_cases import delaunay_board
    monkeypatch.setenv("PADNE_B200_SPMV", spmv)
    for fs in (synth.board_2layer(200, n_via=6, size_mm=30.0, seed=11).to_flat(),
               synth.board_4layer_regulator(100, n_via=4, size_mm=30.0, seed=5).to_flat(),
               delaunay_board(3000, seed=5).to_flat()):
        indptr, indices, data, rhs = gpu_assemble(ctx, fs)
        v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
        assert info.spmv_kind == (1 if spmv == "xs" else 0)
        v_ref, _, _ = oracle.solve(indptr, indices, data, rhs)
        assert info.rel_residual <= RESIDUAL_RTOL and voltage_error(v, v_ref, fs.n_nodes) <= VOLTAGE_RTOL


# ---- BASELINE configs 3 and 4 at full size ---------------------------------------------------------------------
def _source_known_answers(fs, v, tol=1e-9):
    from padne_b200.flat import KIND_REGULATOR, KIND_VOLTAGE
    for k in np.nonzero((fs.e_kind == KIND_VOLTAGE) | (fs.e_kind == KIND_REGULATOR))[0]:
        p, q = fs.e_node[k, 0], fs.e_node[k, 1]
        assert abs(v[p] - v[q] - fs.e_val[k]) < tol, (k, v[p] - v[q], fs.e_val[k])     # solver.py:503-511, 526-529
    assert abs(v[fs.i_gnd]) < tol                                                        # solver.py:558-560


def test_full_size_config3_regulator_4M(ctx):
    # 4 layers, 32 x 32 vias, two REGULATORs (the ldo cancellation + gain 0.52): non-symmetric -> BiCGSTAB
    fs = synth.board_4layer_regulator(1000, n_via=32).to_flat()
    assert fs.N > 3.9e6 and fs.has_regulator
    s = build_stamps(fs)
    rhs = s[4]
    ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *s[:4])
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
    assert info.method == "bicgstab" and info.precond == "two-level"
    assert info.rel_residual <= RESIDUAL_RTOL
    assert info.residual_norm < max(1e-9, 16 * info.floor_abs)      # the reference's bar where fp64 can resolve it
    _source_known_answers(fs, v)
    # regulator currents obey their gain: i_sense = gain * i_out leaves the ground current at the floor
    assert abs(info.ground_node_current) < 1e-6
    v2, info2 = ctx.solve(rhs, n_nodes=fs.n_nodes)
    assert np.array_equal(v, v2) and info.iterations == info2.iterations               # bit-reproducible
    v3, _ = ctx.solve(2.0 * rhs, n_nodes=fs.n_nodes)                                   # linearity
    assert voltage_error(v3, 2.0 * v, fs.n_nodes) <= 1e-8


def test_full_size_config4_8layer_10M(ctx):
    # 8-layer 300 x 300 mm backplane, 64 x 64 vias, 4 voltage sources, 16 sinks: symmetric -> two-level PCG
    fs = synth.board_8layer(1118, n_via=64).to_flat()
    assert fs.N > 9.6e6
    s = build_stamps(fs)
    ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *s[:4])
    v, info = ctx.solve(s[4], n_nodes=fs.n_nodes)
    assert info.method == "pcg" and info.precond == "two-level"
    assert info.rel_residual <= RESIDUAL_RTOL
    assert info.residual_norm < max(1e-9, 16 * info.floor_abs)
    _source_known_answers(fs, v)
    assert abs(info.ground_node_current) < 1e-6
    v2, info2 = ctx.solve(s[4], n_nodes=fs.n_nodes)
    assert np.array_equal(v, v2) and info.iterations == info2.iterations


@pytest.mark.parametrize("name,board", [
    ("regulator_250", lambda: synth.board_4layer_regulator(250, n_via=8)),      # 248k DOF: spsolve 24 s
    ("8layer_200", lambda: synth.board_8layer(200, n_via=8)),                   # 315k DOF: spsolve 40 s
])
def test_large_sub_boards_against_spsolve(ctx, name, board):
    # the largest boards of the config-3 / config-4 families SuperLU finishes in under a minute on one core
    fs = board().to_flat()
    indptr, indices, data, rhs = gpu_assemble(ctx, fs)
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
    v_ref, res_ref, ig_ref = oracle.solve(indptr, indices, data, rhs)
    assert info.rel_residual <= RESIDUAL_RTOL
    assert voltage_error(v, v_ref, fs.n_nodes) <= VOLTAGE_RTOL
    # at this size the reference's own residual is 1e-9 ... 1e-8 (~1.3x the evaluation floor u amax ||u||) and its
    # ground current 4e-8 ... 2e-7 (it would warn); the Krylov loop aims at 4x the floor (pdn_opts.atol), so it
    # has to land within 4x of SuperLU's residual and within its ground current
    assert info.residual_norm <= max(4.0 * res_ref, 1e-9)
    assert info.residual_norm <= 8.0 * info.floor_abs
    assert abs(info.ground_node_current) <= max(4.0 * abs(ig_ref), 1e-8)
