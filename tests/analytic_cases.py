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


def delaunay_board(n_pts=1500, seed=0, n_links=6):
    """Two layers of UNSTRUCTURED triangles -- Delaunay of a jittered point set, so no slivers but vertex valences
    4..10 and irregular orbits like a quality mesher's output (``synth`` only makes valence-6 grids) -- one layer
    with a ragged hole, tied by resistors placed on the highest-valence vertices (rows longer than 10 entries),
    with one voltage and one current source."""
    from scipy.spatial import Delaunay, cKDTree
    rng = np.random.default_rng(seed)
    b = synth.SynthBoard(name=f"delaunay_{n_pts}_{seed}")
    m = max(8, int(round(np.sqrt(n_pts))))
    pitch = 10.0 / (m - 1)
    clouds, valences = [], []
    for layer, hole in enumerate((0.0, 1.6)):
        ii, jj = np.meshgrid(np.arange(m), np.arange(m), indexing="ij")
        jit = (rng.random((m, m, 2)) - 0.5) * 0.8 * pitch
        jit[0, :, 0] = jit[-1, :, 0] = 0.0                                     # the outline stays the square
        jit[:, 0, 1] = jit[:, -1, 1] = 0.0
        pts = (np.stack([ii, jj], -1) * pitch + jit).reshape(-1, 2)
        pts = pts[np.linalg.norm(pts - 5.0, axis=1) >= hole]
        t = Delaunay(pts).simplices.astype(np.int64)
        a, bb, c = pts[t[:, 0]], pts[t[:, 1]], pts[t[:, 2]]
        area2 = (bb[:, 0] - a[:, 0]) * (c[:, 1] - a[:, 1]) - (bb[:, 1] - a[:, 1]) * (c[:, 0] - a[:, 0])
        t[area2 < 0] = t[area2 < 0][:, (0, 2, 1)]                              # all counter-clockwise
        t = t[np.linalg.norm((a + bb + c) / 3 - 5.0, axis=1) >= 0.8 * hole]    # Delaunay fills the hole: reopen it
        used = np.unique(t)                                                    # drop vertices no triangle uses
        remap = np.full(len(pts), -1, np.int64)
        remap[used] = np.arange(used.size)
        t = remap[t]
        b.add_mesh(b.add_layer(f"L{layer}.Cu", 2.0 + layer), pts[used], t.astype(np.int32))
        clouds.append(pts[used])
        valences.append(np.bincount(t.reshape(-1), minlength=used.size))
    off = b.v_off
    trees = [cKDTree(c) for c in clouds]

    def at(layer, x, y):
        return int(off[layer] + trees[layer].query((x, y))[1])

    busiest = np.argsort(-valences[0], kind="stable")[:n_links]               # resistors between the layers
    for k, v0 in enumerate(busiest.tolist()):
        x, y = clouds[0][v0]
        b.add_resistors([int(off[0] + v0)], [at(1, x, y)], 1e-3 * (1 + k))
    b.add_voltage_source(at(0, 0.5, 5.0), at(0, 9.5, 5.0), 1.25)
    b.add_current_source(at(1, 5.0, 9.5), at(1, 5.0, 0.5), 0.7)
    return b
