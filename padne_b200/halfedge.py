"""Row F2 of SURVEY.md section 8f: triangle soup <-> the reference's half-edge ``Mesh`` without the
per-edge Python bookkeeping of ``Mesh.from_triangle_soup`` (``padne/mesh.py:302-378``).

Either side of the hot path the reference converts between two data formats: the mesher's output
lists (``cgal_output['vertices']``, ``cgal_output['triangles']``, ``padne/mesh.py:778-786``) become a
pointer graph of ``Vertex`` / ``HalfEdge`` / ``Face`` objects, and the solver turns that graph back
into indices (``[e.origin.i for e in face.edges]``, ``mesh.py:148-160``).  The kernels need the flat
arrays; ``ui`` / ``paraview`` / pickle need the objects.  This module

* derives the complete half-edge topology of a soup as flat integer arrays (:func:`soup_topology`),
  in exactly the numbering ``from_triangle_soup`` produces: half-edge pairs in first-seen order of
  the directed-edge scan (``connect_vertices``, ``mesh.py:265-294``), ``vertex.out`` = the LAST
  interior half-edge leaving the vertex (``mesh.py:322``), ``face.edge`` = the face's third half-edge
  (``mesh.py:323``), boundary loops by "the boundary half-edge that starts where this one ends"
  (``mesh.py:356-373``), ``ValueError('Non-manifold mesh')`` when a vertex starts two boundary
  half-edges (``mesh.py:340-341``);
* instantiates the reference's own classes from those arrays in bulk (:func:`mesh_from_triangle_soup`),
  remembering the flat arrays of every mesh it built in a weak side table, so that
  :func:`padne_b200.flatten.mesh_to_arrays` never walks that mesh again (nothing is added to the mesh
  objects: the pickled ``.padne`` is unchanged);
* can stand in for ``Mesh.from_triangle_soup`` while the reference's mesher runs
  (:func:`fast_triangle_soup`), which is how ``padne_b200.solver.solve`` uses it.

The one thing the reference leaves to chance is kept deterministic here: it starts every boundary loop
from ``set.pop()`` over identity-hashed half-edges (``mesh.py:346``), so the order of
``mesh.boundaries`` and the ``edge`` each boundary face points at vary from run to run there; here the
loops are numbered by their lowest half-edge index and start at it.  ``next`` / ``prev`` / ``face``
membership are identical.

Soups the array path does not reproduce bit for bit -- a directed edge used by two triangles
(``mesh.py:314-325`` silently re-assigns ``hedge.face``), a triangle that repeats a vertex -- are
handed to the reference's own ``from_triangle_soup``, so every error it raises is raised by it.

Host code, numpy only: this is a data-format row, there is no arithmetic to put on the device.
"""
from __future__ import annotations

import contextlib
import dataclasses
import gc
import operator
import weakref
from dataclasses import dataclass

import numpy as np


@dataclass
class SoupTopology:
    """Half-edge arrays of one triangle soup, in ``from_triangle_soup``'s numbering.

    ``he_twin`` is implicit: half-edges are created in pairs, ``twin(h) = h ^ 1``."""
    n_vertices: int
    tri: np.ndarray            # [T,3] i64, as given
    he_origin: np.ndarray      # [H] origin vertex
    he_next: np.ndarray        # [H]
    he_prev: np.ndarray        # [H]
    he_face: np.ndarray        # [H] interior face index, or -1 - loop for boundary half-edges
    corner_he: np.ndarray      # [T,3] half-edge of corner c -> corner c+1
    v_out: np.ndarray          # [V] -1 for a vertex no triangle uses
    f_edge: np.ndarray         # [T]
    loop_edge: np.ndarray      # [B] first (lowest) half-edge of every boundary loop

    @property
    def n_halfedges(self) -> int:
        return int(self.he_origin.shape[0])


class IrregularSoup(ValueError):
    """The soup needs the reference's own sequential construction (see the module docstring)."""


def soup_topology(n_vertices: int, triangles) -> SoupTopology:
    """Half-edge topology of ``triangles`` over ``n_vertices`` points (``mesh.py:302-378``)."""
    tri = np.ascontiguousarray(np.asarray(triangles, dtype=np.int64).reshape(-1, 3))
    n_t = tri.shape[0]
    nv = int(n_vertices)
    if n_t and (tri.min() < 0 or tri.max() >= nv):
        raise IndexError("list index out of range")                 # vertices[i], mesh.py:312
    u = tri.reshape(-1)                                              # scan order of mesh.py:314-319
    v = tri[:, (1, 2, 0)].reshape(-1)
    if np.any(u == v):
        raise IrregularSoup("triangle repeats a vertex")
    lo, hi = np.minimum(u, v), np.maximum(u, v)
    _, first, inverse = np.unique(lo * nv + hi, return_index=True, return_inverse=True)
    # pair number = rank of the undirected edge's first appearance in the scan (mesh.py:282-285)
    order = np.argsort(first, kind="stable")
    rank = np.empty_like(order)
    rank[order] = np.arange(order.shape[0])
    pair = rank[inverse]
    u_first = u[first][inverse]                                      # direction the pair was created in
    he = 2 * pair + (u != u_first)
    n_h = 2 * order.shape[0]
    if np.unique(he).shape[0] != he.shape[0]:
        raise IrregularSoup("directed edge used by two triangles")
    he_origin = np.empty(n_h, np.int64)
    he_origin[he] = u
    he_origin[he ^ 1] = v
    he_face = np.full(n_h, -1, np.int64)
    he_face[he] = np.arange(3 * n_t) // 3
    corner_he = he.reshape(-1, 3)
    he_next = np.full(n_h, -1, np.int64)
    he_prev = np.full(n_h, -1, np.int64)
    nxt = corner_he[:, (1, 2, 0)].reshape(-1)
    he_next[he] = nxt
    he_prev[nxt] = he
    # vertex.out: the last scan position whose origin is the vertex (mesh.py:322)
    v_out = np.full(nv, -1, np.int64)
    if n_t:
        uniq, last_rev = np.unique(u[::-1], return_index=True)
        v_out[uniq] = he[u.shape[0] - 1 - last_rev]
    f_edge = corner_he[:, 2].copy()
    # boundary half-edges in index order; one per origin or the mesh is not manifold (mesh.py:333-343)
    bnd = np.flatnonzero(he_face < 0)
    b_origin = he_origin[bnd]
    if np.unique(b_origin).shape[0] != b_origin.shape[0]:
        raise ValueError("Non-manifold mesh")
    start_at = np.full(nv, -1, np.int64)
    start_at[b_origin] = bnd
    b_next = start_at[he_origin[bnd ^ 1]]                            # starts where this one ends (mesh.py:362-363)
    he_next[bnd] = b_next
    he_prev[b_next] = bnd
    # loops: the boundary is O(sqrt(N)) half-edges, a plain walk is enough
    loop_edge = []
    nxt_l = he_next.tolist() if bnd.shape[0] else []
    seen = set()
    for h in bnd.tolist():
        if h in seen:
            continue
        code = -1 - len(loop_edge)
        loop_edge.append(h)
        members = []
        while h not in seen:
            seen.add(h)
            members.append(h)
            h = nxt_l[h]
        he_face[members] = code
    return SoupTopology(n_vertices=nv, tri=tri, he_origin=he_origin, he_next=he_next, he_prev=he_prev,
                        he_face=he_face, corner_he=corner_he, v_out=v_out, f_edge=f_edge,
                        loop_edge=np.asarray(loop_edge, np.int64))


def face_arrays(topo: SoupTopology) -> np.ndarray:
    """``[[e.origin.i for e in f.edges] for f in mesh.faces]`` (``mesh.py:148-160``): the walk starts at
    ``face.edge``, the third half-edge, so every triangle comes back rotated to ``(v3, v1, v2)``."""
    return np.ascontiguousarray(topo.tri[:, (2, 0, 1)].astype(np.int32))


# ------------------------------------------------------------------------------------------------
# flat arrays of meshes built here (or registered by a caller that has them), keyed by identity
# ------------------------------------------------------------------------------------------------
_arrays_of = weakref.WeakKeyDictionary()


def register_arrays(mesh, xy, tri) -> None:
    """Remember ``(xy [n,2] f64, tri [m,3] i32 in mesh.faces order)`` for ``mesh``."""
    xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2)
    tri = np.ascontiguousarray(tri, np.int32).reshape(-1, 3)
    xy.setflags(write=False)
    tri.setflags(write=False)
    _arrays_of[mesh] = (xy, tri)


def known_arrays(mesh):
    """The registered arrays of ``mesh`` if they still describe it (same vertex and face counts)."""
    try:
        hit = _arrays_of.get(mesh)
    except TypeError:                                                # unhashable / not weak-referenceable
        return None
    if hit is None or hit[0].shape[0] != len(mesh.vertices) or hit[1].shape[0] != len(mesh.faces):
        return None
    return hit


def _fill_store(store, objects) -> None:
    """``IndexStore.add`` for a whole list (``mesh.py:204-206``); the objects already carry ``i``."""
    if hasattr(store, "_idx_to_obj") and len(store) == 0:
        store._idx_to_obj = objects
    else:
        for o in objects:
            store.add(o)


def mesh_from_triangle_soup(mesh_mod, points, triangles, mesh_cls=None):
    """``Mesh.from_triangle_soup(points, triangles)`` (``padne/mesh.py:302-378``) built from
    :func:`soup_topology`.  ``mesh_mod`` is the reference's ``padne.mesh`` module; ``points`` is a list of
    its ``Point`` objects (kept, not copied, like the reference) or an ``[n,2]`` array."""
    gc_was_on = gc.isenabled()
    gc.disable()    # ~9 live objects per vertex are created here: generational scans of them only cost time
    try:
        return _mesh_from_triangle_soup(mesh_mod, points, triangles, mesh_cls or mesh_mod.Mesh)
    finally:
        if gc_was_on:
            gc.enable()


def _mesh_from_triangle_soup(mesh_mod, points, triangles, mesh_cls):
    original = mesh_cls.__dict__.get("_pdn_original_from_triangle_soup")
    if isinstance(points, np.ndarray):
        xy = np.ascontiguousarray(points, np.float64).reshape(-1, 2)
        Point = mesh_mod.Point
        points = [Point(x, y) for x, y in xy.tolist()]
    else:
        points = list(points)
        xy = np.array([(p.x, p.y) for p in points], np.float64).reshape(-1, 2)
    triangles = list(triangles) if not isinstance(triangles, np.ndarray) else triangles
    if not isinstance(triangles, np.ndarray):
        assert all(len(t) == 3 for t in triangles)                   # mesh.py:311
    try:
        topo = soup_topology(len(points), triangles if len(triangles) else np.zeros((0, 3), np.int64))
    except IrregularSoup:
        build = original.__func__ if original is not None else mesh_cls.from_triangle_soup.__func__
        return build(mesh_cls, points, triangles)
    msh = mesh_cls()
    vertices, hedges, faces, loops = _instantiate(mesh_mod, points, topo)
    _fill_store(msh.vertices, vertices)
    _fill_store(msh.halfedges, hedges)
    _fill_store(msh.faces, faces)
    _fill_store(msh.boundaries, loops)
    if hasattr(msh, "_edge_map"):                                      # mesh.py:288-289
        org = topo.he_origin.tolist()
        msh._edge_map.update(zip(zip(org, _dest(org)), hedges))
    register_arrays(msh, xy, face_arrays(topo))
    return msh


_FIELDS = {"Vertex": ("p", "out", "i"), "HalfEdge": ("origin", "twin", "next", "prev", "face", "i"),
           "Face": ("edge", "is_boundary", "i")}


def _pick(objects, indices):
    """``[objects[i] for i in indices]`` in C."""
    if not indices:
        return []
    if len(indices) == 1:
        return [objects[indices[0]]]
    return list(operator.itemgetter(*indices)(objects))


def _instantiate(mesh_mod, points, topo: SoupTopology):
    """The reference's ``Vertex`` / ``HalfEdge`` / ``Face`` objects of ``topo``, fully linked.

    The objects are plain (non-slot) dataclasses (``mesh.py:72-76, 87-94, 142-146``): each is allocated
    with ``__new__`` and given its complete attribute dictionary at once, instead of a constructor call
    followed by four attribute stores.  Should the reference ever change those fields, the constructors
    are used."""
    Vertex, HalfEdge, Face, index_type = mesh_mod.Vertex, mesh_mod.HalfEdge, mesh_mod.Face, mesh_mod.index_type
    n_v, n_h, n_t, n_l = len(points), topo.n_halfedges, topo.tri.shape[0], topo.loop_edge.shape[0]
    idx = list(np.arange(max(n_v, n_h, n_t, n_l, 1), dtype=index_type))    # the reference's np.uint64 indices
    face_of = np.where(topo.he_face >= 0, topo.he_face, n_t - 1 - topo.he_face).tolist()   # loops follow the faces
    plain = all(tuple(f.name for f in dataclasses.fields(c)) == _FIELDS[c.__name__] and not hasattr(c, "__slots__")
                for c in (Vertex, HalfEdge, Face))
    if not plain:
        vertices = [Vertex(p, None, i) for p, i in zip(points, idx)]
        hedges = [HalfEdge(o, None, None, None, None, i) for o, i in zip(_pick(vertices, topo.he_origin.tolist()), idx)]
        faces = [Face(e, False, i) for e, i in zip(_pick(hedges, topo.f_edge.tolist()), idx)]
        loops = [Face(e, True, i) for e, i in zip(_pick(hedges, topo.loop_edge.tolist()), idx)]
        both = faces + loops
        for h, n, p, f in zip(hedges, _pick(hedges, topo.he_next.tolist()), _pick(hedges, topo.he_prev.tolist()),
                              _pick(both, face_of)):
            h.next, h.prev, h.face = n, p, f
        for a, b in zip(hedges[0::2], hedges[1::2]):
            a.twin, b.twin = b, a
        for vtx, o in zip(vertices, topo.v_out.tolist()):
            if o >= 0:
                vtx.out = hedges[o]
        return vertices, hedges, faces, loops
    vertices = [object.__new__(Vertex) for _ in range(n_v)]
    hedges = [object.__new__(HalfEdge) for _ in range(n_h)]
    faces = [object.__new__(Face) for _ in range(n_t)]
    loops = [object.__new__(Face) for _ in range(n_l)]
    twins = [None] * n_h
    twins[0::2], twins[1::2] = hedges[1::2], hedges[0::2]
    outs = _pick(hedges + [None], topo.v_out.tolist())                 # -1 (no triangle uses the vertex) -> None
    for obj, d in zip(vertices, [{"p": p, "out": o, "i": i} for p, o, i in zip(points, outs, idx)]):
        obj.__dict__ = d
    for obj, d in zip(hedges, [{"origin": o, "twin": t, "next": n, "prev": p, "face": f, "i": i}
                               for o, t, n, p, f, i in zip(_pick(vertices, topo.he_origin.tolist()), twins,
                                                           _pick(hedges, topo.he_next.tolist()),
                                                           _pick(hedges, topo.he_prev.tolist()),
                                                           _pick(faces + loops, face_of), idx)]):
        obj.__dict__ = d
    for obj, d in zip(faces, [{"edge": e, "is_boundary": False, "i": i}
                              for e, i in zip(_pick(hedges, topo.f_edge.tolist()), idx)]):
        obj.__dict__ = d
    for obj, d in zip(loops, [{"edge": e, "is_boundary": True, "i": i}
                              for e, i in zip(_pick(hedges, topo.loop_edge.tolist()), idx)]):
        obj.__dict__ = d
    return vertices, hedges, faces, loops


def _dest(org):
    """Destination of every half-edge = origin of its twin (pairs are adjacent)."""
    dst = [0] * len(org)
    dst[0::2] = org[1::2]
    dst[1::2] = org[0::2]
    return dst


@contextlib.contextmanager
def fast_triangle_soup(mesh_mod):
    """While active, ``mesh_mod.Mesh.from_triangle_soup`` is :func:`mesh_from_triangle_soup` -- the
    mesher's call at ``padne/mesh.py:783-786`` then yields meshes whose flat arrays are already known."""
    cls = mesh_mod.Mesh
    original = cls.__dict__["from_triangle_soup"]
    cls._pdn_original_from_triangle_soup = original
    cls.from_triangle_soup = classmethod(
        lambda c, points, triangles: mesh_from_triangle_soup(mesh_mod, points, triangles, mesh_cls=c))
    try:
        yield
    finally:
        cls.from_triangle_soup = original
        del cls._pdn_original_from_triangle_soup
