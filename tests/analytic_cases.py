"""Problems with a known answer, after the reference's own synthetic / end-to-end tests, built on structured
meshes (the CGAL mesher is not available) and run through ``padne_b200.solver.solve_meshed`` with whatever context
the caller passes: the oracle-backed one on the CPU, the real one on the GPU.  TEST INFRASTRUCTURE ONLY.

* :func:`check_linear_rectangle` -- ``tests/test_solver.py:461-595``: a 10 x 1 rectangle, five 1 V sources from
  points on the left edge to points on the right edge; every source holds its 1 V exactly and the potential is
  linear in x to 0.05 V (the reference's own tolerance).
* :func:`check_superposition` -- ``tests/test_solver.py:1449-1564``: a voltage source and a current source; the
  solution with both equals the sum of the solutions with the other source zeroed (0 V / 0 A).
"""
import numpy as np

import miniref
from padne_b200 import solver as S
from padne_b200 import synth


def _solve(board, ctx):
    prob, meshes, m2l = synth.to_reference_objects(board, miniref, miniref, miniref)
    sol = S.solve_meshed(prob, meshes, m2l, ref=miniref.solver, ctx=ctx)
    return np.concatenate([z.values for ls in sol.layer_solutions for z in ls.potentials]), sol


def check_linear_rectangle(ctx):
    nx, ny, pitch = 101, 11, 0.1
    xy, tri, gmap = synth.grid_layer(nx, ny, pitch, seed=3, jitter=0.15)
    b = synth.SynthBoard(name="linear_rectangle")
    b.add_mesh(b.add_layer("TestLayer", 1.0), xy, tri)
    rows = [1, 3, 5, 7, 9]
    assert gmap.min() >= 0
    for j in rows:
        b.add_voltage_source(int(gmap[nx - 1, j]), int(gmap[0, j]), 1.0)      # positive on the right
    v, sol = _solve(b, ctx)
    assert len(sol.layer_solutions) == 1
    left = np.array([v[gmap[0, j]] for j in rows])
    right = np.array([v[gmap[nx - 1, j]] for j in rows])
    np.testing.assert_allclose(right - left, 1.0, rtol=0, atol=1e-6)
    assert right.mean() > left.mean()
    slope = (right.mean() - left.mean()) / (xy[:, 0].max() - xy[:, 0].min())
    assert np.max(np.abs(v - (left.mean() + (xy[:, 0] - xy[:, 0].min()) * slope))) < 0.05
    assert abs(sol.solver_info.ground_node_current) < 1e-6


def check_superposition(ctx):
    def board(volts, amps):
        b, g, _ = synth.plane_stack("superposition", 2, 40, 20.0, 2, 5, 0.2)
        assert min(g[0][2, 20], g[0][37, 20], g[1][20, 37], g[1][20, 2]) >= 0        # none inside a via hole
        b.add_voltage_source(int(g[0][2, 20]), int(g[0][37, 20]), volts)
        b.add_current_source(int(g[1][20, 37]), int(g[1][20, 2]), amps)
        return b

    v_full, _ = _solve(board(1.5, 0.8), ctx)
    v_volt, _ = _solve(board(1.5, 0.0), ctx)
    v_amp, _ = _solve(board(0.0, 0.8), ctx)
    scale = np.max(np.abs(v_full))
    assert scale > 0.1
    assert np.max(np.abs(v_full - (v_volt + v_amp))) <= 1e-6 * scale
