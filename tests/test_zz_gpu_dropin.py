"""The drop-in flow ``padne_b200.solver.solve_meshed`` on the GPU, through the real ``capi.Context``.

The reference's package is absent on the GPU box, so the objects are the stand-in classes of ``tests/miniref.py``
(checked against the unmodified reference on the CPU by ``tests/test_dropin.py``); the checker is the same flow with
the oracle behind the context (``tests/oracle_context.py``: C assembly restatement + the reference's ``spsolve``
call).  Tolerances are the north star's: voltages within 1e-6 relative, relative residual <= 1e-10, pattern and
values of the assembled matrix bit-exact.  Runs last (file name) because it needs every other piece to work.
"""
import numpy as np
import pytest

import analytic_cases
import miniref
from oracle_context import OracleContext
from padne_b200 import paraview
from padne_b200 import solver as S
from padne_b200 import synth

pytestmark = pytest.mark.gpu

BOARDS = {
    "via_tht_standin": lambda: synth.board_via_tht_standin(20, seed=1, n_via=2),
    "2layer": lambda: synth.board_2layer(60, n_via=4, size_mm=30.0, seed=8),
    "4layer_regulator": lambda: synth.board_4layer_regulator(26, n_via=2, size_mm=12.0, seed=4),
    "8layer": lambda: synth.board_8layer(16, n_via=1, size_mm=8.0, seed=6),
    "delaunay": lambda: analytic_cases.delaunay_board(1500, seed=0),     # unstructured: valences 3..12
}


@pytest.mark.parametrize("name", sorted(BOARDS))
def test_solve_meshed_on_the_gpu_against_the_oracle_flow(ctx, name, tmp_path):
    prob, meshes, m2l = synth.to_reference_objects(BOARDS[name](), miniref, miniref, miniref)
    want = S.solve_meshed(prob, meshes, m2l, ref=miniref.solver, ctx=OracleContext())
    got = S.solve_meshed(prob, meshes, m2l, ref=miniref.solver, ctx=ctx)

    assert type(got) is miniref.Solution and len(got.layer_solutions) == len(prob.layers)
    info = S.last_solve_info
    assert info is not None and info.rel_residual <= 1e-10
    scale = max(np.max(np.abs(z.values)) for ls in want.layer_solutions for z in ls.potentials)
    for la, lb in zip(got.layer_solutions, want.layer_solutions):
        assert [id(m) for m in la.meshes] == [id(m) for m in lb.meshes]
        for za, zb, pa, pb in zip(la.potentials, lb.potentials, la.power_densities, lb.power_densities):
            assert np.max(np.abs(za.values - zb.values)) <= 1e-6 * scale          # north star: 1e-6 relative
            np.testing.assert_allclose(pa.values, pb.values, rtol=1e-4, atol=1e-9 * max(np.max(pb.values), 1e-300))
    # row F4 on the result: the exporter writes one file per layer with the solved voltages
    paraview.export_solution(got, tmp_path)
    assert len(list(tmp_path.iterdir())) == len(set(m2l))


def test_assemble_system_object_form_is_bit_exact(ctx):
    prob, meshes, m2l = synth.to_reference_objects(BOARDS["4layer_regulator"](), miniref, miniref, miniref)
    vindex = miniref.VertexIndexer.create(meshes)
    nets = list(prob.networks)
    nidx = S.node_indexer_create(miniref.solver, prob, meshes, m2l, vindex, nets, ctx)
    nidx_o = S.node_indexer_create(miniref.solver, prob, meshes, m2l, vindex, nets, OracleContext())
    assert list(nidx.node_to_global_index.values()) == list(nidx_o.node_to_global_index.values())
    L, r = S.assemble_system(prob, meshes, m2l, vindex, nets, nidx, ctx)
    Lo, ro = S.assemble_system(prob, meshes, m2l, vindex, nets, nidx, OracleContext())
    assert np.array_equal(L.indptr, Lo.indptr) and np.array_equal(L.indices, Lo.indices)
    assert np.array_equal(L.data, Lo.data) and np.array_equal(r, ro)


def test_linear_rectangle_known_answer_on_the_gpu(ctx):
    analytic_cases.check_linear_rectangle(ctx)              # tests/test_solver.py:461-595 of the reference


def test_superposition_principle_on_the_gpu(ctx):
    analytic_cases.check_superposition(ctx)                 # tests/test_solver.py:1449-1564 of the reference


@pytest.mark.parametrize("n_pts", [1500, 30000])
def test_unstructured_triangles_assembly_bit_exact_and_solve(ctx, n_pts):
    """Delaunay clouds instead of the valence-6 grids of ``synth``: vertex orbits of 3..12 triangles, ragged
    boundaries, a hole, rows longer than the SpMV's shared-memory slice width (its direct-load path)."""
    import scipy.sparse
    from oracle import oracle
    from padne_b200.stamps import build_stamps
    fs = analytic_cases.delaunay_board(n_pts, seed=n_pts).to_flat()
    s_row, s_col, s_val, s_op, rhs = build_stamps(fs)
    ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, s_row, s_col, s_val, s_op)
    indptr, indices, data = ctx.get_csr()
    o_indptr, o_indices, o_data = oracle.assemble(fs, (s_row, s_col, s_val, s_op))
    assert np.array_equal(indptr, o_indptr) and np.array_equal(indices, o_indices)
    assert np.array_equal(data, o_data)                                            # values bit-exact
    assert np.diff(indptr)[:fs.n_nodes].max() > 10                                 # wider than a TMA slice slot
    v, info = ctx.solve(rhs, n_nodes=fs.n_nodes)
    v_ref, _, _ = oracle.solve(o_indptr, o_indices, o_data, rhs)                   # the reference's spsolve call
    assert info.rel_residual <= 1e-10
    scale = np.max(np.abs(v_ref[:fs.n_nodes]))
    assert np.max(np.abs(v[:fs.n_nodes] - v_ref[:fs.n_nodes])) <= 1e-6 * scale
    L = scipy.sparse.csr_matrix((data, indices, indptr), shape=(fs.N, fs.N))
    assert np.linalg.norm(L @ v - rhs) <= 1e-8 * max(1.0, np.linalg.norm(rhs))
