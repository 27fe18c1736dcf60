"""Golden fixtures for row F2 (half-edge topology of a triangle soup), read off the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_halfedge.py

Every array is an attribute walk over the ``Mesh`` that the reference's own ``Mesh.from_triangle_soup``
(padne/mesh.py:302-378) built from the stored soup: ``he_origin[i] = halfedges[i].origin.i`` and so on;
``he_face`` is the interior face index or -1 (boundary); ``face_vertices`` is
``[[e.origin.i for e in f.edges] for f in mesh.faces]`` (padne/mesh.py:148-160).  The reference numbers
its boundary faces from ``set.pop()`` over identity-hashed objects (padne/mesh.py:346), so the loops are
stored as the lowest half-edge index of each (``loop_min``); membership is pinned through ``he_next``.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.refload import load_reference  # noqa: E402
from padne_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
problem, mesh, solver = load_reference()


def halfedge_case(name, xy, tri):
    pts = [mesh.Point(float(x), float(y)) for x, y in xy]
    msh = mesh.Mesh.from_triangle_soup(pts, [tuple(int(i) for i in t) for t in tri])
    hs = list(msh.halfedges)
    assert all(int(h.twin.i) == int(h.i) ^ 1 for h in hs)
    loop_min = [min(int(e.i) for e in f.edges) for f in msh.boundaries]
    np.savez_compressed(
        os.path.join(HERE, f"halfedge_{name}.npz"), n_vertices=len(pts), tri=np.asarray(tri, np.int64).reshape(-1, 3),
        he_origin=np.array([int(h.origin.i) for h in hs], np.int64),
        he_next=np.array([int(h.next.i) for h in hs], np.int64),
        he_prev=np.array([int(h.prev.i) for h in hs], np.int64),
        he_face=np.array([-1 if h.face.is_boundary else int(h.face.i) for h in hs], np.int64),
        v_out=np.array([-1 if v.out is None else int(v.out.i) for v in msh.vertices], np.int64),
        f_edge=np.array([int(f.edge.i) for f in msh.faces], np.int64),
        face_vertices=np.array([[int(e.origin.i) for e in f.edges] for f in msh.faces], np.int32).reshape(-1, 3),
        loop_min=np.array(sorted(loop_min), np.int64))
    print(f"halfedge_{name}: {len(pts)} vertices, {len(hs)} half-edges, {len(msh.faces)} faces, {len(loop_min)} loops")


if __name__ == "__main__":
    from test_halfedge import _soups
    for name, xy, tri in _soups():
        if name in ("hole", "complex", "two_islands_and_a_loose_vertex", "grid", "grid_via_holes", "grid_shuffled"):
            halfedge_case(name, xy, tri)
