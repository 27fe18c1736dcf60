"""Reference object graph -> :class:`~padne_b200.flat.FlatSystem`.

Duck-typed on the reference's own classes so that the caller keeps using ``padne.mesh`` /
``padne.problem`` objects untouched:

* a mesh is anything with ``.vertices`` (items having ``.p.x``, ``.p.y``) and ``.faces`` (items whose
  ``.edges`` yield half-edges with ``.origin.i``) -- ``padne/mesh.py:148-160, 203-205`` -- or any
  object that already carries arrays as ``.xy`` / ``.tri`` (the fast path that skips the O(N)
  Python walk, row F2 of SURVEY.md section 8f);
* ``vindex`` / ``node_indexer`` are the reference's ``VertexIndexer`` / ``NodeIndexer`` instances
  (``padne/solver.py:216-229, 350-466``), consumed, never recomputed;
* elements are matched by class name exactly like the ``match`` statement of
  ``padne/solver.py:474-541``; anything else raises the same ``NotImplementedError``.
"""
from __future__ import annotations

import operator

import numpy as np

from . import halfedge
from .flat import (FlatSystem, KIND_CURRENT, KIND_REGULATOR, KIND_RESISTOR, KIND_VOLTAGE)


def mesh_to_arrays(msh):
    """``(xy [n,2] f64, tri [m,3] i32)`` of one mesh, triangles in ``mesh.faces`` order.

    Three sources, cheapest first: arrays the object carries (``.xy`` / ``.tri``), arrays remembered
    for this mesh by :mod:`padne_b200.halfedge` (meshes built by ``mesh_from_triangle_soup``, or walked
    once already), and the walk over the reference's object graph, which is then remembered."""
    if hasattr(msh, "xy") and hasattr(msh, "tri"):
        return (np.ascontiguousarray(msh.xy, np.float64).reshape(-1, 2),
                np.ascontiguousarray(msh.tri, np.int32).reshape(-1, 3))
    hit = halfedge.known_arrays(msh)
    if hit is not None:
        return hit
    xy = np.array([(p.x, p.y) for p in [v.p for v in msh.vertices]], dtype=np.float64).reshape(-1, 2)
    tri = _walk_faces(msh)
    try:
        halfedge.register_arrays(msh, xy, tri)
    except TypeError:                      # a duck-typed mesh that cannot be weakly referenced
        pass
    return xy, tri


def _walk_faces(msh) -> np.ndarray:
    """``[[e.origin.i for e in f.edges] for f in mesh.faces]`` (``padne/mesh.py:148-160``).  On the reference's
    half-edge objects the three corners are read by pointer (``face.edge``, ``.next``, ``.next.next``) instead of
    through the ``edges`` generator; any face that is not a closed triangle takes the generic walk."""
    try:
        nxt, corner = operator.attrgetter("next"), operator.attrgetter("origin.i")
        e0 = [f.edge for f in msh.faces]
        e1 = list(map(nxt, e0))
        e2 = list(map(nxt, e1))
        if all(map(operator.is_, map(nxt, e2), e0)):
            n = len(e0)
            return np.stack([np.fromiter(map(corner, e), np.int64, n) for e in (e0, e1, e2)],
                            axis=1).astype(np.int32).reshape(-1, 3)
    except AttributeError:                 # a duck-typed mesh that only offers ``face.edges``
        pass
    return np.array([[int(e.origin.i) for e in f.edges] for f in msh.faces], dtype=np.int32).reshape(-1, 3)


def find_best_ground_node_index(prob, node_indexer) -> int:
    """``padne/solver.py:671-686``: ``n`` of the VoltageSource with the largest voltage, scanning
    ``prob.networks`` (not the filtered list) and exact VoltageSource instances only."""
    max_voltage = float("-inf")
    ground_node_index = 0
    for network in prob.networks:
        for element in network.elements:
            if type(element).__name__ != "VoltageSource":
                continue
            if element.voltage > max_voltage:
                max_voltage = element.voltage
                ground_node_index = node_indexer.node_to_global_index[element.n]
    return int(ground_node_index)


def flatten_elements(filtered_networks, node_indexer):
    """Element table in stamping order (``padne/solver.py:808-809`` then ``:473``)."""
    n2g = node_indexer.node_to_global_index
    x2g = node_indexer.extra_source_to_global_index
    kind, node, val, gain, xidx = [], [], [], [], []
    for network in filtered_networks:
        for e in network.elements:
            name = type(e).__name__
            if name == "Resistor":
                kind.append(KIND_RESISTOR); node.append((n2g[e.a], n2g[e.b], 0, 0))
                val.append(e.resistance); gain.append(0.0); xidx.append(-1)
            elif name == "CurrentSource":
                kind.append(KIND_CURRENT); node.append((n2g[e.f], n2g[e.t], 0, 0))
                val.append(e.current); gain.append(0.0); xidx.append(-1)
            elif name == "VoltageSource":
                kind.append(KIND_VOLTAGE); node.append((n2g[e.p], n2g[e.n], 0, 0))
                val.append(e.voltage); gain.append(0.0); xidx.append(x2g[e])
            elif name == "VoltageRegulator":
                kind.append(KIND_REGULATOR)
                node.append((n2g[e.v_p], n2g[e.v_n], n2g[e.s_f], n2g[e.s_t]))
                val.append(e.voltage); gain.append(e.gain); xidx.append(x2g[e])
            else:
                raise NotImplementedError(f"Unsupported node type {e}")
    return (np.array(kind, np.int32), np.array(node, np.int32).reshape(-1, 4),
            np.array(val, np.float64), np.array(gain, np.float64), np.array(xidx, np.int32))


def flatten_system(prob, meshes, mesh_index_to_layer_index, vindex, filtered_networks,
                   node_indexer, name: str = "") -> FlatSystem:
    """Everything ``assemble_system`` (``padne/solver.py:783-812``) reads, as arrays."""
    arrays = [mesh_to_arrays(m) for m in meshes]
    nv = [a[0].shape[0] for a in arrays]
    nt = [a[1].shape[0] for a in arrays]
    v_off = np.concatenate([[0], np.cumsum(nv)]).astype(np.int64)
    t_off = np.concatenate([[0], np.cumsum(nt)]).astype(np.int64)
    if vindex is not None and len(vindex.global_index_to_vertex_index) != int(v_off[-1]):
        raise ValueError("VertexIndexer does not match the meshes")
    xy = np.concatenate([a[0] for a in arrays]) if arrays else np.zeros((0, 2))
    tri = np.concatenate([a[1] for a in arrays]) if arrays else np.zeros((0, 3), np.int32)
    cond = np.array([prob.layers[mesh_index_to_layer_index[i]].conductance
                     for i in range(len(meshes))], np.float64)
    kind, node, val, gain, xidx = flatten_elements(filtered_networks, node_indexer)
    return FlatSystem(v_off=v_off, t_off=t_off, xy=xy, tri=tri, conductance=cond,
                      n_internal=int(node_indexer.internal_node_count),
                      n_extra=len(node_indexer.extra_source_to_global_index),
                      i_gnd=find_best_ground_node_index(prob, node_indexer),
                      e_kind=kind, e_node=node, e_val=val, e_gain=gain, e_xidx=xidx, name=name)
