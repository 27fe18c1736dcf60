"""Row F2: ``padne_b200.halfedge`` against the reference's own
``Mesh.from_triangle_soup`` (``padne/mesh.py:302-378``) -- CPU only.

The known-answer cases restate ``tests/test_mesh.py:527-810`` of the reference (single triangle, two
triangles, star, the non-manifold pair, the star-like complex mesh, the square with a hole); the
comparisons against the unmodified reference need ``/root/reference`` (build container only) and are
skipped elsewhere; ``tests/golden/halfedge_*.npz`` carries the reference's answers to the GPU box.
"""
import glob
import os
import pickle

import numpy as np
import pytest

from conftest import GOLDEN
from padne_b200 import halfedge, synth
from padne_b200.flatten import mesh_to_arrays


def _reference():
    from oracle import refload
    if not refload.reference_available():
        pytest.skip("reference not present")
    return refload.load_reference()


def _soups():
    """(name, xy, tri) of the cases: the reference's own test meshes + synthetic layers."""
    out = [
        ("single", [(0, 0), (1, 0), (0, 1)], [(0, 1, 2)]),
        ("two", [(0, 0), (1, 0), (1, 1), (0, 1)], [(0, 1, 2), (0, 2, 3)]),
        ("star", [(0, 0), (0, 1), (-1, 0), (0, -1), (1, 0)], [(0, 4, 1), (0, 1, 2), (0, 2, 3), (0, 3, 4)]),
        ("complex", [(-1, -1), (1, -1), (1, 1), (-1, 1), (-2, 0), (0, -2), (2, 0), (0, 2)],
         [(0, 1, 2), (0, 2, 3), (0, 3, 4), (1, 0, 5), (2, 1, 6), (3, 2, 7)]),
        ("hole", [(0, 0), (4, 0), (4, 4), (0, 4), (1, 1), (3, 1), (3, 3), (1, 3)],
         [(0, 1, 4), (1, 5, 4), (1, 2, 5), (2, 6, 5), (2, 3, 6), (3, 7, 6), (3, 0, 7), (0, 4, 7)]),
        ("two_islands_and_a_loose_vertex", [(0, 0), (1, 0), (0, 1), (5, 5), (6, 5), (5, 6), (9, 9)],
         [(3, 4, 5), (0, 1, 2)]),
        ("empty", [], []),
    ]
    out = [(n, np.array(p, np.float64).reshape(-1, 2), np.array(t, np.int64).reshape(-1, 3)) for n, p, t in out]
    xy, tri, _ = synth.grid_layer(13, 9, 0.5, seed=3)
    out.append(("grid", xy, tri.astype(np.int64)))
    xy, tri, _ = synth.grid_layer(24, 20, 0.25, seed=5, via_blocks=[(3, 3), (12, 9)])
    out.append(("grid_via_holes", xy, tri.astype(np.int64)))
    rng = np.random.default_rng(11)                      # triangle order and rotation must not matter
    perm = rng.permutation(tri.shape[0])
    rot = np.array([np.roll(t, int(k)) for t, k in zip(tri[perm], rng.integers(0, 3, tri.shape[0]))])
    out.append(("grid_shuffled", xy, rot.astype(np.int64)))
    # unstructured clouds (what a quality mesher emits: valences 3..10, ragged boundary), one with a hole
    from scipy.spatial import Delaunay
    for name, n_pts, hole in (("delaunay_400", 400, 0.0), ("delaunay_annulus_900", 900, 0.28)):
        pts = np.random.default_rng(n_pts).random((n_pts, 2))
        t = Delaunay(pts).simplices.astype(np.int64)
        a, b, c = pts[t[:, 0]], pts[t[:, 1]], pts[t[:, 2]]
        flip = (b[:, 0] - a[:, 0]) * (c[:, 1] - a[:, 1]) - (b[:, 1] - a[:, 1]) * (c[:, 0] - a[:, 0]) < 0
        t[flip] = t[flip][:, (0, 2, 1)]                                   # all counter-clockwise
        keep = np.linalg.norm((a + b + c) / 3 - 0.5, axis=1) >= hole
        out.append((name, pts, t[keep]))
    return out


def _describe(msh):
    """Everything index-valued about a Mesh, as plain ints (boundary faces as canonical loops)."""
    hs = list(msh.halfedges)
    loops = set()
    for f in msh.boundaries:
        cyc = [int(e.i) for e in f.edges]
        k = cyc.index(min(cyc))
        loops.add(tuple(cyc[k:] + cyc[:k]))
    return dict(
        n=(len(msh.vertices), len(msh.halfedges), len(msh.faces), len(msh.boundaries)),
        points=[(v.p.x, v.p.y) for v in msh.vertices],
        v_i=[int(v.i) for v in msh.vertices],
        v_out=[-1 if v.out is None else int(v.out.i) for v in msh.vertices],
        he=[(int(h.i), int(h.origin.i), int(h.twin.i), int(h.next.i), int(h.prev.i),
             bool(h.face.is_boundary), -1 if h.face.is_boundary else int(h.face.i)) for h in hs],
        f_edge=[int(f.edge.i) for f in msh.faces],
        f_i=[int(f.i) for f in msh.faces],
        loops=loops,
        edge_map={(int(a), int(b)): int(h.i) for (a, b), h in msh._edge_map.items()},
    )


@pytest.mark.parametrize("case", _soups(), ids=lambda c: c[0])
def test_same_object_graph_as_the_reference(case):
    _, mesh, ref = _reference()
    _, xy, tri = case
    pts = [mesh.Point(float(x), float(y)) for x, y in xy]
    tris = [tuple(int(i) for i in t) for t in tri]
    stock = mesh.Mesh.from_triangle_soup(pts, tris)
    ours = halfedge.mesh_from_triangle_soup(mesh, pts, tris)
    assert type(ours) is mesh.Mesh
    assert _describe(ours) == _describe(stock)
    assert all(a.p is b for a, b in zip(ours.vertices, pts))             # the caller's Point objects are kept
    assert all(type(v.i) is mesh.index_type for v in ours.vertices)
    assert all(type(h.i) is mesh.index_type for h in ours.halfedges)
    # the flat arrays it remembers are what a walk over the stock mesh yields (mesh.py:148-160)
    sxy = np.array([(v.p.x, v.p.y) for v in stock.vertices], np.float64).reshape(-1, 2)
    stri = np.array([[int(e.origin.i) for e in f.edges] for f in stock.faces], np.int32).reshape(-1, 3)
    kxy, ktri = halfedge.known_arrays(ours)
    assert np.array_equal(kxy, sxy) and np.array_equal(ktri, stri)
    assert mesh_to_arrays(ours)[1] is ktri                                # no second walk
    # the reference's own operator reads the two graphs identically, emission order included
    if len(tris) and all(v.out is not None for v in stock.vertices):     # a loose vertex breaks orbit() there too
        a, b = ref.laplace_operator(ours), ref.laplace_operator(stock)
        assert np.array_equal(a.row, b.row) and np.array_equal(a.col, b.col) and np.array_equal(a.data, b.data)
    # pickle round trip (the .padne format, mesh.py:238-259)
    back = pickle.loads(pickle.dumps(ours))
    assert _describe(back) == _describe(stock)
    assert sorted(vars(ours)) == sorted(vars(stock))                       # nothing extra travels in the pickle


def test_reference_errors_are_the_reference_s():
    _, mesh, _ = _reference()
    pts = [mesh.Point(0.0, 0.0), mesh.Point(1.0, 0.0), mesh.Point(0.0, 1.0), mesh.Point(0.0, -1.0)]
    with pytest.raises(ValueError):                                        # tests/test_mesh.py:713-733
        halfedge.mesh_from_triangle_soup(mesh, pts, [(0, 1, 2), (0, 1, 3)])
    # two fans touching in one vertex: two boundary half-edges start there (mesh.py:340-341)
    bow = [mesh.Point(0.0, 0.0), mesh.Point(1.0, 0.0), mesh.Point(1.0, 1.0), mesh.Point(-1.0, 0.0), mesh.Point(-1.0, -1.0)]
    for build in (mesh.Mesh.from_triangle_soup, lambda p, t: halfedge.mesh_from_triangle_soup(mesh, p, t)):
        with pytest.raises(ValueError, match="Non-manifold mesh"):
            build(bow, [(0, 1, 2), (0, 3, 4)])
    with pytest.raises(IndexError):
        halfedge.mesh_from_triangle_soup(mesh, pts, [(0, 1, 7)])


def test_fast_triangle_soup_swaps_and_restores():
    _, mesh, _ = _reference()
    stock_fn = mesh.Mesh.__dict__["from_triangle_soup"]
    pts = [mesh.Point(0.0, 0.0), mesh.Point(1.0, 0.0), mesh.Point(0.0, 1.0)]
    with halfedge.fast_triangle_soup(mesh):
        m = mesh.Mesh.from_triangle_soup(pts, [(0, 1, 2)])                 # the call of mesh.py:783
        assert halfedge.known_arrays(m) is not None
        with pytest.raises(ValueError):                                    # irregular soups reach the stock builder
            mesh.Mesh.from_triangle_soup(pts + [mesh.Point(0.0, -1.0)], [(0, 1, 2), (0, 1, 3)])
    assert mesh.Mesh.__dict__["from_triangle_soup"] is stock_fn
    assert not hasattr(mesh.Mesh, "_pdn_original_from_triangle_soup")
    assert halfedge.known_arrays(mesh.Mesh.from_triangle_soup(pts, [(0, 1, 2)])) is None


# ---- portable part: known answers and the committed reference topology ---------------------------
def test_known_answers_of_the_reference_tests():
    # tests/test_mesh.py:527-810: (vertices, faces, half-edges, boundaries)
    expect = {"single": (3, 1, 6, 1), "two": (4, 2, 10, 1), "star": (5, 4, 16, 1),
              "complex": (8, 6, 26, 1), "hole": (8, 8, 32, 2)}
    for name, xy, tri in _soups():
        if name not in expect:
            continue
        t = halfedge.soup_topology(xy.shape[0], tri)
        assert (t.n_vertices, tri.shape[0], t.n_halfedges, t.loop_edge.shape[0]) == expect[name]
        assert t.n_vertices - t.n_halfedges // 2 + tri.shape[0] == (0 if name == "hole" else 1)   # Euler
        # structure (assert_mesh_structure_valid of the reference's tests): next/prev inverse, faces closed
        assert np.array_equal(t.he_prev[t.he_next], np.arange(t.n_halfedges))
        h = np.arange(t.n_halfedges)
        assert np.array_equal(t.he_origin[t.he_next], t.he_origin[h ^ 1])      # next starts where this one ends
        interior = t.he_face >= 0
        assert np.array_equal(t.he_next[t.he_next[t.he_next[h[interior]]]], h[interior])
        assert np.array_equal(t.he_face[t.he_next], t.he_face)
    two = halfedge.soup_topology(4, [(0, 1, 2), (0, 2, 3)])
    assert len(_walk(two, int(two.loop_edge[0]))) == 4                         # tests/test_mesh.py:603-604


def _walk(t, h0):
    out, h = [], h0
    while True:
        out.append(h)
        h = int(t.he_next[h])
        if h == h0:
            return out


def test_non_manifold_and_irregular_soups():
    with pytest.raises(ValueError, match="Non-manifold mesh"):
        halfedge.soup_topology(5, [(0, 1, 2), (0, 3, 4)])
    with pytest.raises(halfedge.IrregularSoup):
        halfedge.soup_topology(4, [(0, 1, 2), (0, 1, 3)])
    with pytest.raises(halfedge.IrregularSoup):
        halfedge.soup_topology(3, [(0, 1, 1)])
    with pytest.raises(IndexError):
        halfedge.soup_topology(3, [(0, 1, 3)])


def golden_halfedge_files():
    return sorted(glob.glob(os.path.join(GOLDEN, "halfedge_*.npz")))


@pytest.mark.parametrize("path", golden_halfedge_files(), ids=os.path.basename)
def test_topology_golden(path):
    # arrays read off the UNMODIFIED reference's Mesh.from_triangle_soup (tests/golden/make_golden_halfedge.py)
    d = np.load(path)
    t = halfedge.soup_topology(int(d["n_vertices"]), d["tri"])
    assert np.array_equal(t.he_origin, d["he_origin"])
    assert np.array_equal(t.he_next, d["he_next"])
    assert np.array_equal(t.he_prev, d["he_prev"])
    assert np.array_equal(np.where(t.he_face >= 0, t.he_face, -1), d["he_face"])
    assert np.array_equal(t.v_out, d["v_out"])
    assert np.array_equal(t.f_edge, d["f_edge"])
    assert np.array_equal(halfedge.face_arrays(t), d["face_vertices"])
    ours = sorted(min(c) for c in (_walk(t, int(h)) for h in t.loop_edge))
    assert ours == sorted(int(x) for x in d["loop_min"])


def test_golden_files_exist():
    assert len(golden_halfedge_files()) >= 4
