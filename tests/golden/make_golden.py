"""Generate the golden fixtures in this directory by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

For every case the reference's own code path is executed on reference objects:
``Mesh.from_triangle_soup`` (padne/mesh.py:302) -> ``VertexIndexer.create`` (solver.py:222) ->
``NodeIndexer.create`` (solver.py:399) -> ``assemble_system`` (solver.py:783) ->
``solve_system`` (solver.py:767) -> ``produce_layer_solutions`` (solver.py:578).
Stored per case: the flat inputs (``FlatSystem`` arrays, with the reference's own node numbering),
``L.tocsr()`` (indptr / indices / data), ``r``, ``v``, ``SolverInfo`` and the per-face power
densities.  Mesh-only cases store ``laplace_operator(mesh)`` (solver.py:171) as emitted COO.
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.refload import load_reference  # noqa: E402
from padne_b200 import synth  # noqa: E402
from padne_b200.flatten import flatten_system  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
problem, mesh, solver = load_reference()
import shapely.geometry as geometry  # noqa: E402  (the loader's stub)


def laplace_case(name, xy, tri):
    pts = [mesh.Point(float(x), float(y)) for x, y in xy]
    msh = mesh.Mesh.from_triangle_soup(pts, [tuple(int(i) for i in t) for t in tri])
    L = solver.laplace_operator(msh)
    csr = L.tocsr()
    csr.sort_indices()
    np.savez_compressed(os.path.join(HERE, f"laplace_{name}.npz"),
                        xy=np.asarray(xy, np.float64), tri=np.asarray(tri, np.int32),
                        coo_row=L.row.astype(np.int64), coo_col=L.col.astype(np.int64), coo_data=L.data,
                        indptr=csr.indptr.astype(np.int32), indices=csr.indices.astype(np.int32),
                        data=csr.data)
    print(f"laplace_{name}: n={len(pts)} nnz={csr.nnz}")


def system_case(name, board):
    prob, meshes, m2l = synth.to_reference_objects(board, problem, mesh, geometry)
    vindex = solver.VertexIndexer.create(meshes)
    filtered = list(prob.networks)
    nidx = solver.NodeIndexer.create(prob, meshes, m2l, vindex, filtered)
    L, r = solver.assemble_system(prob, meshes, m2l, vindex, filtered, nidx)
    with warnings.catch_warnings():
        warnings.simplefilter("error")          # a singular-matrix warning must not go unnoticed
        v, info = solver.solve_system(L, r)
    csr = L.tocsr()
    csr.sort_indices()
    assert csr.indices.dtype == np.int32
    layer_solutions = solver.produce_layer_solutions(prob.layers, vindex, meshes, m2l, v,
                                                     [[] for _ in prob.layers])
    pd = np.concatenate([p.values for ls in layer_solutions for p in ls.power_densities])
    fs = flatten_system(prob, meshes, m2l, vindex, filtered, nidx, name=name)
    d = fs.to_npz_dict()
    d.update(indptr=csr.indptr.astype(np.int32), indices=csr.indices.astype(np.int32), data=csr.data,
             r=r, v=v, residual_norm=np.float64(info.residual_norm),
             ground_node_current=np.float64(info.ground_node_current), power_density=pd,
             mesh_layer=np.asarray(m2l, np.int32),
             layer_conductance=np.array([l.conductance for l in prob.layers]))
    np.savez_compressed(os.path.join(HERE, f"system_{name}.npz"), **d)
    print(f"system_{name}: N={fs.N} nnz={csr.nnz} residual={info.residual_norm:.3e} "
          f"ground_current={info.ground_node_current:.3e}")


def main():
    # reference known answer: unit square + centre (tests/test_solver.py:776-852)
    laplace_case("unit_square", [(0, 0), (1, 0), (1, 1), (0, 1), (0.5, 0.5)],
                 [(0, 1, 4), (1, 2, 4), (2, 3, 4), (3, 0, 4)])
    xy, tri, _ = synth.grid_layer(12, 9, 0.37, seed=3, jitter=0.2)
    laplace_case("jitter_12x9", xy, tri)
    xy, tri, _ = synth.grid_layer(9, 9, 0.5, seed=0, jitter=0.0)           # exact right angles
    laplace_case("right_angle_9x9", xy, tri)
    xy, tri, _ = synth.grid_layer(16, 14, 0.25, seed=5, jitter=0.2, via_blocks=[(3, 4), (9, 6)])
    laplace_case("holes_16x14", xy, tri)
    rng = np.random.default_rng(11)                                         # obtuse-heavy: |cot| path
    xy, tri, _ = synth.grid_layer(10, 10, 1.0, seed=7, jitter=0.45)
    laplace_case("obtuse_10x10", xy * np.array([1.0, 0.23]), tri[rng.permutation(len(tri))])

    system_case("via_tht_standin_20", synth.board_via_tht_standin(20, seed=1, n_via=2))
    system_case("2layer_24", synth.board_2layer(24, n_via=2, size_mm=10.0, seed=2))
    system_case("2layer_right_angle_20", synth.board_2layer(20, n_via=2, size_mm=9.5, seed=0, jitter=0.0))
    system_case("4layer_regulator_26", synth.board_4layer_regulator(26, n_via=2, size_mm=12.0, seed=4))
    system_case("8layer_16", synth.board_8layer(16, n_via=1, size_mm=8.0, seed=6))
    system_case("2layer_60", synth.board_2layer(60, n_via=4, size_mm=30.0, seed=8))


if __name__ == "__main__":
    main()
