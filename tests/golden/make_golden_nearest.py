"""Golden fixtures for row F3 (nearest mesh vertex), produced by the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_nearest.py

The KD-trees are built by the reference's own ``NodeIndexer._construct_kdtrees``
(padne/solver.py:356-396) on reference ``Mesh`` objects; every query is the reference's statement pair
``_, i = kdtree.query((x, y), k=1)``; ``layer_global_index_and_vertex[layer][i][0]`` (padne/solver.py:421-424).
Query points: every Connection of the board (they sit on vertices) plus random points of the board area
(generic position, so no two vertices are equidistant).  Stored: the flat inputs of ``pdn_nearest_vertex``
and the global vertex index the reference returns.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.refload import load_reference  # noqa: E402
from padne_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
problem, mesh, solver = load_reference()
import shapely.geometry as geometry  # noqa: E402  (the loader's stub)


def nearest_case(name, board, n_random, seed):
    prob, meshes, m2l = synth.to_reference_objects(board, problem, mesh, geometry)
    vindex = solver.VertexIndexer.create(meshes)
    trees, index_and_vertex = solver.NodeIndexer._construct_kdtrees(prob, meshes, m2l, vindex)
    xy = np.array([(v.p.x, v.p.y) for m in meshes for v in m.vertices], np.float64)
    v_off = np.concatenate([[0], np.cumsum([len(m.vertices) for m in meshes])]).astype(np.int64)
    q_xy, q_layer = [], []
    for net in prob.networks:
        for conn in net.connections:
            q_xy.append((conn.point.x, conn.point.y))
            q_layer.append(prob.layers.index(conn.layer))
    rng = np.random.default_rng(seed)
    lo, hi = xy.min(0) - 0.3, xy.max(0) + 0.3           # some queries fall outside the copper
    for _ in range(n_random):
        q_xy.append(tuple(rng.uniform(lo, hi)))
        q_layer.append(int(rng.integers(len(prob.layers))))
    expect = []
    for (x, y), layer in zip(q_xy, q_layer):
        _, i = trees[layer].query((x, y), k=1)
        expect.append(index_and_vertex[layer][i][0])
    np.savez_compressed(os.path.join(HERE, f"nearest_{name}.npz"), v_off=v_off, mesh_layer=np.asarray(m2l, np.int32),
                        xy=xy, q_xy=np.asarray(q_xy, np.float64), q_layer=np.asarray(q_layer, np.int32),
                        expect=np.asarray(expect, np.int64))
    print(f"nearest_{name}: {len(meshes)} meshes, {xy.shape[0]} vertices, {len(expect)} queries")


if __name__ == "__main__":
    nearest_case("2layer_24", synth.board_2layer(24, n_via=2, size_mm=10.0, seed=2), 400, 1)
    nearest_case("8layer_16", synth.board_8layer(16, n_via=1, size_mm=8.0, seed=6), 600, 2)
    nearest_case("via_tht_standin_20", synth.board_via_tht_standin(20, seed=1, n_via=2), 300, 3)
