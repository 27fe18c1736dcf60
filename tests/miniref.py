"""A few dozen lines of stand-in for the reference's ``padne.problem`` / ``padne.mesh`` / ``padne.solver`` CLASSES,
so that the drop-in flow ``padne_b200.solver.solve_meshed`` can run where the reference is absent (the GPU box).
TEST INFRASTRUCTURE ONLY; written for these tests, no arithmetic of the hot path lives here.

Only what ``solve_meshed`` touches exists: the data model of a problem (``padne/problem.py:22-212``, names and field
names as the product matches them by name), meshes that carry their flat arrays, ``ZeroForm`` / ``TwoForm`` value
holders (``padne/mesh.py:381-387, 575-584``), the two indexers (``padne/solver.py:216-229, 350-466``) and the result
classes (``solver.py:33-52``).  ``tests/test_dropin.py`` checks on the CPU that a board run through these classes gets
the voltages the unmodified reference computes for it.  Internal nodes are numbered in first-seen order here (the
reference: ``set`` iteration order), which permutes unknowns beyond the vertices and nothing else.

Use: ``prob, meshes, m2l = synth.to_reference_objects(board, miniref, miniref, miniref)``; ``ref=miniref.solver``.
"""
import types
from dataclasses import dataclass, field

import numpy as np


# ---- geometry / problem ------------------------------------------------------------------------------
class Point:
    def __init__(self, x, y):
        self.x, self.y = float(x), float(y)


class MultiPolygon:
    def __init__(self, geoms=()):
        self.geoms = tuple(geoms)


class NodeID:
    pass


@dataclass(frozen=True, eq=False)
class Layer:
    shape: object
    name: str
    conductance: float


@dataclass(frozen=True, eq=False)
class Connection:
    layer: Layer
    point: Point
    node_id: NodeID = field(default_factory=NodeID)


@dataclass(frozen=True, eq=False)
class Resistor:
    a: NodeID
    b: NodeID
    resistance: float
    extra_variable_count = 0
    terminals = property(lambda self: [self.a, self.b])


@dataclass(frozen=True, eq=False)
class CurrentSource:
    f: NodeID
    t: NodeID
    current: float
    extra_variable_count = 0
    terminals = property(lambda self: [self.f, self.t])


@dataclass(frozen=True, eq=False)
class VoltageSource:
    p: NodeID
    n: NodeID
    voltage: float
    extra_variable_count = 1
    terminals = property(lambda self: [self.p, self.n])


@dataclass(frozen=True, eq=False)
class VoltageRegulator:
    v_p: NodeID
    v_n: NodeID
    s_f: NodeID
    s_t: NodeID
    voltage: float
    gain: float
    extra_variable_count = 1
    terminals = property(lambda self: [self.v_p, self.v_n, self.s_f, self.s_t])


class Network:
    def __init__(self, connections, elements):
        self.connections, self.elements = list(connections), list(elements)
        self.nodes = {}
        for e in self.elements:
            for t in e.terminals:
                self.nodes.setdefault(t, len(self.nodes))


@dataclass
class Problem:
    layers: list
    networks: list
    project_name: str = ""


# ---- meshes and forms --------------------------------------------------------------------------------
class Mesh:
    """Carries ``xy`` / ``tri`` (what ``flatten.mesh_to_arrays`` looks for first); ``vertices`` / ``faces`` only
    have a length."""

    @classmethod
    def from_triangle_soup(cls, points, triangles):
        m = cls()
        m.xy = np.array([(p.x, p.y) for p in points], np.float64).reshape(-1, 2)
        m.tri = np.array(list(triangles), np.int32).reshape(-1, 3)
        m.vertices, m.faces = range(len(m.xy)), range(len(m.tri))
        return m


class ZeroForm:
    def __init__(self, mesh):
        self.mesh, self.values = mesh, np.zeros(len(mesh.vertices))


class TwoForm:
    def __init__(self, mesh):
        self.mesh, self.values = mesh, np.zeros(len(mesh.faces))


# ---- indexers and results ------------------------------------------------------------------------------
@dataclass
class VertexIndexer:
    global_index_to_vertex_index: list

    @classmethod
    def create(cls, meshes):
        return cls([(mi, vi) for mi, m in enumerate(meshes) for vi in range(len(m.vertices))])


class _BruteTree:
    def __init__(self, xy):
        self.xy = xy

    def query(self, point, k=1):
        d = (self.xy[:, 0] - point[0]) ** 2 + (self.xy[:, 1] - point[1]) ** 2
        i = int(np.argmin(d))
        return float(d[i]) ** 0.5, i


@dataclass
class NodeIndexer:
    node_to_global_index: dict
    extra_source_to_global_index: dict
    internal_node_count: int

    @classmethod
    def _construct_kdtrees(cls, prob, meshes, mesh_index_to_layer_index, vindex):
        trees, tables, off = {}, {}, 0
        for m, layer_i in zip(meshes, mesh_index_to_layer_index):
            tables.setdefault(layer_i, []).extend((off + i, None) for i in range(len(m.vertices)))
            off += len(m.vertices)
        all_xy = np.concatenate([m.xy for m in meshes]) if meshes else np.zeros((0, 2))
        for layer_i, table in tables.items():
            if table:
                trees[layer_i] = _BruteTree(all_xy[[g for g, _ in table]])
        return trees, tables

    @classmethod
    def create(cls, prob, meshes, mesh_index_to_layer_index, vindex, filtered_networks):
        trees, tables = cls._construct_kdtrees(prob, meshes, mesh_index_to_layer_index, vindex)
        n2g = {}
        for net in filtered_networks:
            for c in net.connections:
                layer_i = prob.layers.index(c.layer)
                _, k = trees[layer_i].query((c.point.x, c.point.y), k=1)
                g = tables[layer_i][k][0]
                if n2g.setdefault(c.node_id, g) != g:
                    raise ValueError("Duplicate connection vertices found, this should not happen.")
        at = len(vindex.global_index_to_vertex_index)
        n_internal = 0
        for net in filtered_networks:
            for node in net.nodes:
                if node not in n2g:
                    n2g[node] = at
                    at += 1
                    n_internal += 1
        x2g = {}
        for net in filtered_networks:
            for e in net.elements:
                if e.extra_variable_count > 1:
                    raise NotImplementedError("Extra variable count > 1 not supported yet")
                if e.extra_variable_count == 1:
                    x2g[e] = at
                    at += 1
        return cls(n2g, x2g, n_internal)


@dataclass
class SolverInfo:
    ground_node_current: float
    residual_norm: float


@dataclass
class LayerSolution:
    meshes: list
    potentials: list
    power_densities: list
    disconnected_meshes: list


@dataclass
class Solution:
    problem: Problem
    layer_solutions: list
    solver_info: SolverInfo


class SolverWarning(UserWarning):
    pass


mesh = types.SimpleNamespace(Point=Point, Mesh=Mesh, ZeroForm=ZeroForm, TwoForm=TwoForm)
solver = types.SimpleNamespace(VertexIndexer=VertexIndexer, NodeIndexer=NodeIndexer, SolverInfo=SolverInfo,
                               LayerSolution=LayerSolution, Solution=Solution, SolverWarning=SolverWarning, mesh=mesh)
