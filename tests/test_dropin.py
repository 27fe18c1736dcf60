"""The drop-in entry ``padne_b200.solver.solve_meshed`` (steps 4-10 of ``solve``, ``padne/solver.py:860-902``)
end to end on the reference's own objects -- CPU only, needs ``/root/reference`` (build container).

The arithmetic behind the context is the oracle (``tests/oracle_context.py``); what is under test is the host side:
node indexing through the batched lookup, flattening, stamp order, RHS, the classes of the returned ``Solution``,
the ground-current warning and that the resident matrix is not uploaded a second time.  The same flow on the GPU
(real ``capi.Context``, no reference needed) is ``tests/test_zz_gpu_dropin.py``.
"""
import pickle
import warnings

import numpy as np
import pytest

from oracle import refload
from oracle_context import OracleContext
from padne_b200 import solver as S
from padne_b200 import synth

pytestmark = pytest.mark.skipif(not refload.reference_available(), reason="reference not present")

BOARDS = {
    "via_tht_standin": lambda: synth.board_via_tht_standin(18, seed=1, n_via=2),
    "2layer": lambda: synth.board_2layer(16, n_via=2, size_mm=8.0, seed=2),
    "4layer_regulator": lambda: synth.board_4layer_regulator(18, n_via=2, size_mm=8.0, seed=4),
    "8layer": lambda: synth.board_8layer(12, n_via=1, size_mm=6.0, seed=6),
    "delaunay": lambda: __import__("analytic_cases").delaunay_board(700, seed=1),      # unstructured triangles
}


def _reference_solution(prob, meshes, m2l, solver):
    """Steps 4-10 of the unmodified ``solve`` (``padne/solver.py:860-902``)."""
    vindex = solver.VertexIndexer.create(meshes)
    filtered = list(prob.networks)
    nidx = solver.NodeIndexer.create(prob, meshes, m2l, vindex, filtered)
    L, r = solver.assemble_system(prob, meshes, m2l, vindex, filtered, nidx)
    v, info = solver.solve_system(L, r)
    ls = solver.produce_layer_solutions(prob.layers, vindex, meshes, m2l, v, [[] for _ in prob.layers])
    return solver.Solution(problem=prob, layer_solutions=ls, solver_info=info), L, r, v


@pytest.mark.parametrize("name", sorted(BOARDS))
def test_solve_meshed_matches_the_reference_pipeline(name):
    problem, mesh, solver = refload.load_reference()
    import shapely.geometry as geometry
    prob, meshes, m2l = synth.to_reference_objects(BOARDS[name](), problem, mesh, geometry)
    want, L, r, v = _reference_solution(prob, meshes, m2l, solver)

    ctx = OracleContext()
    got = S.solve_meshed(prob, meshes, m2l, ref=solver, ctx=ctx)

    assert type(got) is solver.Solution and got.problem is prob
    assert ctx.calls.count("assemble") == 1 and "set_csr" not in ctx.calls        # the resident matrix is reused
    assert ctx.calls.count("nearest_vertex") == 1 and ctx.calls.count("solve") == 1
    assert type(got.solver_info) is solver.SolverInfo
    assert got.solver_info.residual_norm < 1e-9                                  # tests/test_solver.py:2083-2089
    assert len(got.layer_solutions) == len(prob.layers)
    for a, b in zip(got.layer_solutions, want.layer_solutions):
        assert type(a) is solver.LayerSolution
        assert [id(m) for m in a.meshes] == [id(m) for m in b.meshes] and a.disconnected_meshes == []
        for za, zb, pa, pb in zip(a.potentials, b.potentials, a.power_densities, b.power_densities):
            assert type(za) is mesh.ZeroForm and type(pa) is mesh.TwoForm and za.mesh is zb.mesh and pa.mesh is zb.mesh
            # same matrix, same SuperLU call: the voltages are the reference's to rounding of the solve
            np.testing.assert_allclose(za.values, zb.values, rtol=1e-9, atol=1e-12)
            np.testing.assert_allclose(pa.values, pb.values, rtol=1e-6, atol=1e-18)
    pickle.loads(pickle.dumps(got))                                               # the .padne format, cli.py:223-224


@pytest.mark.parametrize("name", ["4layer_regulator", "delaunay"])
def test_assemble_system_equals_the_reference_matrix(name):
    problem, mesh, solver = refload.load_reference()
    import shapely.geometry as geometry
    prob, meshes, m2l = synth.to_reference_objects(BOARDS[name](), problem, mesh, geometry)
    vindex = solver.VertexIndexer.create(meshes)
    filtered = list(prob.networks)
    nidx = solver.NodeIndexer.create(prob, meshes, m2l, vindex, filtered)
    L_ref, r_ref = solver.assemble_system(prob, meshes, m2l, vindex, filtered, nidx)
    L, r = S.assemble_system(prob, meshes, m2l, vindex, filtered, nidx, OracleContext())
    want = L_ref.tocsr()
    want.sort_indices()
    assert L.shape == L_ref.shape and L.nnz == L_ref.nnz and L.tocsc().shape == L_ref.shape   # benchmarks.py:652-654
    assert np.array_equal(L.indptr, want.indptr) and np.array_equal(L.indices, want.indices)
    assert np.array_equal(L.data, want.data) and np.array_equal(r, r_ref)
    # a matrix that is NOT resident (a plain scipy matrix) is uploaded before the solve
    ctx = OracleContext()
    v, info = S.solve_system(L_ref, r_ref, ctx)
    assert ctx.calls == ["set_csr", "solve"] and type(info) is solver.SolverInfo
    v_ref, _ = solver.solve_system(L_ref, r_ref)
    np.testing.assert_allclose(v, v_ref, rtol=1e-9, atol=1e-12)


def test_ground_current_warning_and_singular_system():
    problem, mesh, solver = refload.load_reference()
    import shapely.geometry as geometry
    board = synth.board_2layer(10, n_via=1, size_mm=5.0, seed=3)
    prob, meshes, m2l = synth.to_reference_objects(board, problem, mesh, geometry)

    class Leaky(OracleContext):
        def solve(self, rhs, n_nodes=-1, **opts):
            v, info = super().solve(rhs, n_nodes, **opts)
            info.ground_node_current = 1e-3
            return v, info

    with pytest.warns(solver.SolverWarning, match="Ground node current is not zero"):     # padne/solver.py:880-888
        S.solve_meshed(prob, meshes, m2l, ref=solver, ctx=Leaky())
    with warnings.catch_warnings():
        warnings.simplefilter("error", solver.SolverWarning)
        S.solve_meshed(prob, meshes, m2l, ref=solver, ctx=OracleContext())                # no warning when it is ~0

    class Singular(OracleContext):
        def solve(self, rhs, n_nodes=-1, **opts):
            from padne_b200 import capi
            raise capi.SingularSystemError("constraint loop")

    from scipy.sparse.linalg import MatrixRankWarning
    with pytest.warns(MatrixRankWarning):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore", solver.SolverWarning)
            sol = S.solve_meshed(prob, meshes, m2l, ref=solver, ctx=Singular())
    assert all(np.isnan(z.values).all() for ls in sol.layer_solutions for z in ls.potentials)   # SuperLU: all-NaN v


# ---- the stand-in classes the GPU box uses (tests/miniref.py) give the reference's voltages ---------------------
@pytest.mark.parametrize("name", sorted(BOARDS))
def test_miniref_flow_gives_the_reference_voltages(name):
    import miniref
    problem, mesh, solver = refload.load_reference()
    import shapely.geometry as geometry
    prob, meshes, m2l = synth.to_reference_objects(BOARDS[name](), problem, mesh, geometry)
    want, _, _, _ = _reference_solution(prob, meshes, m2l, solver)

    mprob, mmeshes, mm2l = synth.to_reference_objects(BOARDS[name](), miniref, miniref, miniref)
    # the stand-in's own NodeIndexer (brute-force trees) and the batched-lookup drop-in agree
    vindex = miniref.VertexIndexer.create(mmeshes)
    a = miniref.NodeIndexer.create(mprob, mmeshes, mm2l, vindex, list(mprob.networks))
    b = S.node_indexer_create(miniref.solver, mprob, mmeshes, mm2l, vindex, list(mprob.networks), OracleContext())
    assert list(a.node_to_global_index.values()) == list(b.node_to_global_index.values())
    assert list(a.extra_source_to_global_index.values()) == list(b.extra_source_to_global_index.values())

    got = S.solve_meshed(mprob, mmeshes, mm2l, ref=miniref.solver, ctx=OracleContext())
    assert type(got) is miniref.Solution and got.solver_info.residual_norm < 1e-9
    for la, lb in zip(got.layer_solutions, want.layer_solutions):
        assert len(la.potentials) == len(lb.potentials)
        for za, zb, pa, pb in zip(la.potentials, lb.potentials, la.power_densities, lb.power_densities):
            np.testing.assert_allclose(za.values, zb.values, rtol=1e-9, atol=1e-12)
            np.testing.assert_allclose(pa.values, pb.values, rtol=1e-6, atol=1e-18)


def test_laplace_operator_and_power_density_seams_on_reference_meshes():
    """``laplace_operator(mesh)`` (``padne/solver.py:171``) and ``compute_power_density`` (``:728``) called the way
    the reference's tests and benchmarks call them, on the reference's ``Mesh`` / ``ZeroForm``."""
    problem, mesh, solver = refload.load_reference()
    xy, tri, _ = synth.grid_layer(9, 7, 0.4, seed=5, jitter=0.25)
    msh = mesh.Mesh.from_triangle_soup([mesh.Point(float(x), float(y)) for x, y in xy],
                                       [tuple(int(i) for i in t) for t in tri])
    want = solver.laplace_operator(msh).tocsr()
    want.sort_indices()
    got = S.laplace_operator(msh, OracleContext())
    assert got.shape == want.shape and type(got).__name__ == "coo_matrix"
    g = got.tocsr()
    g.sort_indices()
    assert np.array_equal(g.indptr, want.indptr) and np.array_equal(g.indices, want.indices)
    assert np.array_equal(g.data, want.data)
    zf = mesh.ZeroForm(msh)
    zf.values[:] = np.random.default_rng(0).random(len(xy))
    pd_want = solver.compute_power_density(zf, 2.0)
    pd_got = S.compute_power_density(zf, 2.0, OracleContext())
    assert type(pd_got) is mesh.TwoForm and pd_got.mesh is msh
    np.testing.assert_allclose(pd_got.values, pd_want.values, rtol=1e-12, atol=0)


# ---- known answers (no reference needed: stand-in classes + the oracle behind the context) ---------------------
def test_linear_rectangle_known_answer():
    import analytic_cases
    analytic_cases.check_linear_rectangle(OracleContext())


def test_superposition_principle():
    import analytic_cases
    analytic_cases.check_superposition(OracleContext())
