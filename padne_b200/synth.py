"""Synthetic multi-layer plane boards as triangle soups.

The shipped KiCad projects need pcbnew + pygerber + shapely + CGAL, none of which exists on the
build or GPU image, so ``BASELINE.json``'s configs are generated directly:

* per layer an ``nx x ny`` structured grid at pitch ``h``; every interior vertex is jittered
  uniformly by ``+-0.2 h`` (``numpy.random.default_rng(seed)``), each cell is split into two
  CCW triangles with alternating diagonal -> about 6.97 stored matrix entries per row, like a
  CGAL mesh;  ``jitter=0`` keeps exact right angles (hypotenuse weights are exactly 0, the
  ``ratio == 0`` skip of ``padne/solver.py:187-190``);
* a via = the 3x3 interior vertices of a 4x4-cell block removed on every layer it spans (a
  square hole with 16 perimeter vertices) and 16 ring resistors per adjacent layer pair, each
  ``16 * R_total`` with the hollow-cylinder ``R_total`` of ``padne/kicad.py:818-836``
  (``kicad.py:1551``);
* layer conductance ``0.035 mm * 5.95e4 S/mm`` (``kicad.py:79, 114-116``).

A board is produced in two equivalent forms: a :class:`SynthBoard` (element terminals as
global vertex ids / internal-node ids) that converts straight to a ``FlatSystem``, and -- for
small sizes, in the build container only -- reference ``problem.Problem`` + ``mesh.Mesh``
objects (``to_reference_objects``) so the real reference can be run on the identical input.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

from .flat import (FlatSystem, KIND_CURRENT, KIND_REGULATOR, KIND_RESISTOR, KIND_VOLTAGE)

COPPER_CONDUCTIVITY = 5.95e4      # S/mm   (kicad.py:79)
COPPER_THICKNESS = 0.035          # mm
LAYER_CONDUCTANCE = COPPER_THICKNESS * COPPER_CONDUCTIVITY
VIA_DRILL = 0.3                   # mm
VIA_PLATING = 0.035               # mm
VIA_RING_POINTS = 16


def via_total_resistance(length_mm: float, drill: float = VIA_DRILL,
                         plating: float = VIA_PLATING,
                         conductivity: float = COPPER_CONDUCTIVITY) -> float:
    """Hollow-cylinder via model, same expression order as ``kicad.py:829-836``."""
    outer_radius = drill / 2 + plating
    inner_radius = drill / 2
    cross_sectional_area = math.pi * (outer_radius ** 2 - inner_radius ** 2)
    return length_mm / (conductivity * cross_sectional_area)


# --------------------------------------------------------------------------------------------
# one layer
# --------------------------------------------------------------------------------------------
def grid_layer(nx: int, ny: int, pitch: float, seed: int = 0, jitter: float = 0.2,
               via_blocks: list[tuple[int, int]] | None = None):
    """One jittered-grid layer.

    ``via_blocks`` lists the lower-left grid vertex ``(i0, j0)`` of each 4x4-cell via block.
    Returns ``(xy [n,2] f64, tri [m,3] i32, gmap [nx,ny] i32)`` where ``gmap[i, j]`` is the
    local vertex index of grid vertex ``(i, j)`` or ``-1`` if a via hole removed it.
    """
    via_blocks = via_blocks or []
    ii, jj = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    x = ii.astype(np.float64) * pitch
    y = jj.astype(np.float64) * pitch
    removed = np.zeros((nx, ny), bool)
    fixed = np.zeros((nx, ny), bool)
    fixed[0, :] = fixed[-1, :] = fixed[:, 0] = fixed[:, -1] = True
    for (i0, j0) in via_blocks:
        if i0 < 1 or j0 < 1 or i0 + 5 > nx - 1 or j0 + 5 > ny - 1:
            raise ValueError("via block too close to the board edge")
        removed[i0 + 1:i0 + 4, j0 + 1:j0 + 4] = True
        fixed[i0:i0 + 5, j0:j0 + 5] = True          # hole perimeter is a boundary: not jittered
    if jitter:
        rng = np.random.default_rng(seed)
        dx = rng.uniform(-jitter * pitch, jitter * pitch, size=(nx, ny))
        dy = rng.uniform(-jitter * pitch, jitter * pitch, size=(nx, ny))
        x = np.where(fixed, x, x + dx)
        y = np.where(fixed, y, y + dy)
    gmap = np.full((nx, ny), -1, np.int32)
    keep = ~removed
    gmap[keep] = np.arange(int(keep.sum()), dtype=np.int32)
    xy = np.stack([x[keep], y[keep]], axis=1)

    # cells -> two CCW triangles, alternating diagonal
    ci, cj = np.meshgrid(np.arange(nx - 1), np.arange(ny - 1), indexing="ij")
    ci, cj = ci.ravel(), cj.ravel()
    v00, v10 = gmap[ci, cj], gmap[ci + 1, cj]
    v11, v01 = gmap[ci + 1, cj + 1], gmap[ci, cj + 1]
    even = ((ci + cj) % 2) == 0
    t1 = np.where(even[:, None], np.stack([v00, v10, v11], 1), np.stack([v00, v10, v01], 1))
    t2 = np.where(even[:, None], np.stack([v00, v11, v01], 1), np.stack([v10, v11, v01], 1))
    tri = np.empty((2 * ci.size, 3), np.int32)
    tri[0::2], tri[1::2] = t1, t2
    tri = tri[(tri >= 0).all(axis=1)]
    return xy, np.ascontiguousarray(tri), gmap


def via_ring(i0: int, j0: int) -> list[tuple[int, int]]:
    """The 16 perimeter grid vertices of the via block at ``(i0, j0)``, counter-clockwise."""
    ring = [(i0 + k, j0) for k in range(4)]
    ring += [(i0 + 4, j0 + k) for k in range(4)]
    ring += [(i0 + 4 - k, j0 + 4) for k in range(4)]
    ring += [(i0, j0 + 4 - k) for k in range(4)]
    return ring


# --------------------------------------------------------------------------------------------
# a board = meshes + networks of lumped elements
# --------------------------------------------------------------------------------------------
@dataclass
class SynthBoard:
    """Meshes plus lumped elements with terminals as global ids.

    Terminal ids ``< n_vertices`` are mesh vertices (global, back-to-back numbering like
    ``VertexIndexer``); ids ``>= n_vertices`` are internal network nodes in creation order.
    Every ``add_*`` call appends ONE network (the reference stamps network by network).
    """
    name: str
    meshes: list = field(default_factory=list)           # [(xy, tri)]
    mesh_layer: list = field(default_factory=list)       # layer index of each mesh
    layer_conductance: list = field(default_factory=list)
    layer_names: list = field(default_factory=list)
    n_internal: int = 0
    # per network: dict(kind=, node=[k,4], val=[k], gain=[k])
    networks: list = field(default_factory=list)

    # -- geometry -----------------------------------------------------------------------
    def add_layer(self, name: str, conductance: float = LAYER_CONDUCTANCE) -> int:
        self.layer_names.append(name)
        self.layer_conductance.append(float(conductance))
        return len(self.layer_names) - 1

    def add_mesh(self, layer: int, xy: np.ndarray, tri: np.ndarray) -> int:
        if self.networks:
            raise RuntimeError("add all meshes before the first network")
        self.meshes.append((np.ascontiguousarray(xy, np.float64), np.ascontiguousarray(tri, np.int32)))
        self.mesh_layer.append(int(layer))
        return len(self.meshes) - 1

    @property
    def v_off(self) -> np.ndarray:
        return np.concatenate([[0], np.cumsum([m[0].shape[0] for m in self.meshes])]).astype(np.int64)

    @property
    def t_off(self) -> np.ndarray:
        return np.concatenate([[0], np.cumsum([m[1].shape[0] for m in self.meshes])]).astype(np.int64)

    @property
    def n_vertices(self) -> int:
        return int(self.v_off[-1])

    def vid(self, mesh_i: int, local) -> np.ndarray:
        """Global vertex id(s) of local vertex index/indices of mesh ``mesh_i``."""
        return np.asarray(local, np.int64) + int(self.v_off[mesh_i])

    def new_internal_node(self) -> int:
        self.n_internal += 1
        return self.n_vertices + self.n_internal - 1

    # -- networks -----------------------------------------------------------------------
    def _add(self, kind, nodes, val, gain=None):
        nodes = np.asarray(nodes, np.int64).reshape(-1, 4)
        k = nodes.shape[0]
        self.networks.append(dict(
            kind=np.full(k, kind, np.int32) if np.isscalar(kind) else np.asarray(kind, np.int32),
            node=nodes, val=np.asarray(val, np.float64).reshape(k),
            gain=np.zeros(k) if gain is None else np.asarray(gain, np.float64).reshape(k)))

    def add_resistors(self, a, b, resistance) -> None:
        a = np.atleast_1d(np.asarray(a, np.int64))
        b = np.atleast_1d(np.asarray(b, np.int64))
        r = np.broadcast_to(np.asarray(resistance, np.float64), a.shape)
        z = np.zeros_like(a)
        self._add(KIND_RESISTOR, np.stack([a, b, z, z], 1), r)

    def add_voltage_source(self, p: int, n: int, voltage: float) -> None:
        self._add(KIND_VOLTAGE, [[p, n, 0, 0]], [voltage])

    def add_current_source(self, f: int, t: int, current: float) -> None:
        self._add(KIND_CURRENT, [[f, t, 0, 0]], [current])

    def add_regulator(self, v_p: int, v_n: int, s_f: int, s_t: int, voltage: float, gain: float) -> None:
        self._add(KIND_REGULATOR, [[v_p, v_n, s_f, s_t]], [voltage], [gain])

    def add_mixed_network(self, kinds, nodes, vals, gains=None) -> None:
        """One network holding several element kinds (e.g. a multi-pad star with 0 V glue)."""
        self._add(np.asarray(kinds, np.int32), nodes, vals, gains)

    # -- flat form ------------------------------------------------------------------------
    def to_flat(self) -> FlatSystem:
        """``FlatSystem`` with this board's own internal-node numbering.

        Extra unknowns follow ``NodeIndexer.create`` (``solver.py:448-460``): one per
        VoltageSource / VoltageRegulator in network/element order.  Ground follows
        ``find_best_ground_node_index`` (``solver.py:671-686``): the ``n`` terminal of the
        first VoltageSource with the strictly largest voltage, else index 0.
        """
        if self.networks:
            kind = np.concatenate([n["kind"] for n in self.networks])
            node = np.concatenate([n["node"] for n in self.networks])
            val = np.concatenate([n["val"] for n in self.networks])
            gain = np.concatenate([n["gain"] for n in self.networks])
        else:
            kind, node = np.zeros(0, np.int32), np.zeros((0, 4), np.int64)
            val = gain = np.zeros(0)
        n_nodes = self.n_vertices + self.n_internal
        has_x = (kind == KIND_VOLTAGE) | (kind == KIND_REGULATOR)
        xidx = np.where(has_x, n_nodes + np.cumsum(has_x) - 1, -1).astype(np.int32)
        i_gnd = 0
        vs = np.nonzero(kind == KIND_VOLTAGE)[0]
        if vs.size:
            i_gnd = int(node[vs[int(np.argmax(val[vs]))], 1])   # argmax = first strict maximum
        return FlatSystem(
            v_off=self.v_off, t_off=self.t_off,
            xy=np.concatenate([m[0] for m in self.meshes]),
            tri=np.concatenate([m[1] for m in self.meshes]),
            conductance=np.array([self.layer_conductance[l] for l in self.mesh_layer]),
            n_internal=self.n_internal, n_extra=int(has_x.sum()), i_gnd=i_gnd,
            e_kind=kind, e_node=node.astype(np.int32), e_val=val, e_gain=gain, e_xidx=xidx,
            name=self.name)


# --------------------------------------------------------------------------------------------
# the BASELINE.json configurations, all size-parametrised
# --------------------------------------------------------------------------------------------
_DIELECTRIC_MM = (0.1, 1.24, 0.1)     # via_tht_4layer stack-up, cycled for deeper stacks


def via_lattice(n: int, k: int) -> list[tuple[int, int]]:
    """Lower-left grid vertices of a ``k x k`` lattice of 4x4-cell via blocks in an ``n x n`` grid."""
    if k <= 0:
        return []
    lo, hi = 2, n - 8
    if hi < lo or (k > 1 and (hi - lo) // (k - 1) < 6):
        raise ValueError(f"{k}x{k} via lattice does not fit a {n}x{n} grid")
    pos = [lo + (hi - lo) // 2] if k == 1 else [lo + (hi - lo) * a // (k - 1) for a in range(k)]
    return [(i, j) for i in pos for j in pos]


def plane_stack(name: str, n_layers: int, n: int, size_mm: float, n_via: int,
                seed: int = 0, jitter: float = 0.2):
    """``n_layers`` planes of ``n x n`` vertices stitched by an ``n_via x n_via`` via lattice.

    Returns ``(board, gmaps, pitch)``; ``gmaps[l][i, j]`` is the GLOBAL id of grid vertex (i, j)
    on layer ``l`` (or -1 inside a via hole).
    """
    b = SynthBoard(name)
    pitch = size_mm / (n - 1)
    blocks = via_lattice(n, n_via)
    gmaps = []
    for l in range(n_layers):
        li = b.add_layer(f"L{l}.Cu")
        xy, tri, gmap = grid_layer(n, n, pitch, seed=seed + l, jitter=jitter, via_blocks=blocks)
        b.add_mesh(li, xy, tri)
        gmaps.append(gmap)
    voff = b.v_off
    gmaps = [np.where(g >= 0, g.astype(np.int64) + int(voff[l]), -1) for l, g in enumerate(gmaps)]
    # one Network of 16 ring resistors per via per adjacent layer pair (kicad.py:1497-1585)
    for (i0, j0) in blocks:
        ring = via_ring(i0, j0)
        ri = np.array([p[0] for p in ring])
        rj = np.array([p[1] for p in ring])
        for l in range(n_layers - 1):
            r_total = via_total_resistance(_DIELECTRIC_MM[l % len(_DIELECTRIC_MM)])
            b.add_resistors(gmaps[l][ri, rj], gmaps[l + 1][ri, rj], r_total * VIA_RING_POINTS)
    return b, gmaps, pitch


def _pick(gmap: np.ndarray, i: int, j: int) -> int:
    """Global id of grid vertex (i, j), or of the nearest vertex to its right that no via hole removed."""
    n = gmap.shape[0]
    for d in range(n):
        for ii in (i + d, i - d):
            if 0 <= ii < n and gmap[ii, j] >= 0:
                return int(gmap[ii, j])
    raise ValueError("no vertex left in grid row")


def board_2layer(n: int = 708, n_via: int = 16, size_mm: float = 100.0, seed: int = 0,
                 jitter: float = 0.2) -> SynthBoard:
    """BASELINE config 2: 2-layer 100x100 mm plane, one VOLTAGE + one CURRENT source (~1M DOF at n=708)."""
    b, g, _ = plane_stack(f"synthetic_2layer_{n}", 2, n, size_mm, n_via, seed, jitter)
    m = n // 2
    b.add_voltage_source(_pick(g[0], 0, m), _pick(g[0], n - 1, m), 1.0)
    b.add_current_source(_pick(g[1], m, n - 1), _pick(g[1], m, 0), 1.0)
    return b


def board_4layer_regulator(n: int = 1000, n_via: int = 32, size_mm: float = 100.0, seed: int = 0,
                           jitter: float = 0.2) -> SynthBoard:
    """BASELINE config 3: 4 layers, via arrays, REGULATOR coupling (non-symmetric), ~4M DOF at n=1000.

    Regulator 1 is the ``ldo`` shape (``tests/kicad/ldo``: ``n=U1.1 f=U1.1 gain=1.0``): ``s_f`` and
    ``v_n`` are the same vertex, so the ``-1`` and ``+gain`` stamps cancel exactly and the entry
    leaves the pattern.  Regulator 2 has four distinct terminals and gain 0.52 (``README.md:180``).
    """
    b, g, _ = plane_stack(f"synthetic_4layer_reg_{n}", 4, n, size_mm, n_via, seed, jitter)
    q, m, t = n // 4, n // 2, (3 * n) // 4
    b.add_voltage_source(_pick(g[0], 1, m), _pick(g[0], n - 2, m), 12.0)
    b.add_regulator(_pick(g[3], q, q), _pick(g[3], q, t), _pick(g[3], q, t), _pick(g[0], t, t), 3.3, 1.0)
    b.add_regulator(_pick(g[2], t, q), _pick(g[2], m, m + 1), _pick(g[1], q, m), _pick(g[1], t, m + 1), 1.8, 0.52)
    b.add_current_source(_pick(g[3], m, 1), _pick(g[3], m, n - 2), 2.0)
    b.add_current_source(_pick(g[1], 1, 1), _pick(g[2], n - 2, n - 2), 0.5)
    return b


def board_8layer(n: int = 1118, n_via: int = 64, size_mm: float = 300.0, seed: int = 0,
                 jitter: float = 0.2, n_layers: int = 8) -> SynthBoard:
    """BASELINE config 4: 8-layer 300x300 mm backplane, 4 voltage sources, 16 current sinks (~10M DOF)."""
    b, g, _ = plane_stack(f"synthetic_{n_layers}layer_{n}", n_layers, n, size_mm, n_via, seed, jitter)
    e = n - 2
    corners = [(1, 1), (e, 1), (e, e), (1, e)]
    for s, (ci, cj) in enumerate(corners):           # 4 sources, top layer corner -> bottom layer corner
        b.add_voltage_source(_pick(g[0], ci, cj), _pick(g[n_layers - 1], *corners[(s + 2) % 4]), 1.0 + 0.25 * s)
    for s in range(16):                              # 16 sinks on a 4x4 raster, alternating layers
        i = (1 + 2 * (s % 4)) * n // 8
        j = (1 + 2 * (s // 4)) * n // 8
        la, lb = s % n_layers, (s + n_layers // 2) % n_layers
        b.add_current_source(_pick(g[la], i, j), _pick(g[lb], i + 1, j + 1), 0.1 * (1 + s % 3))
    return b


def board_via_tht_standin(n: int = 64, seed: int = 0, jitter: float = 0.2, n_via: int = 3) -> SynthBoard:
    """BASELINE config 1 stand-in (``tests/kicad/via_tht_4layer`` cannot be loaded on this image).

    Same topology class: 4 copper layers, via-style stacks, three 0.1 ohm resistors and one 1 V
    source; the resistors hang off pad-like internal nodes glued to copper by 1 mOhm star
    resistors (``kicad.py:535-556, 444``).
    """
    b, g, _ = plane_stack(f"via_tht_4layer_standin_{n}", 4, n, 20.0, n_via, seed, jitter)
    q, t = n // 4, (3 * n) // 4
    b.add_voltage_source(_pick(g[0], q, q), _pick(g[3], q, q), 1.0)
    pads = [((0, t, t), (3, t, t)), ((0, q, t), (1, q, t)), ((2, t, q), (3, t, q))]
    for (la, ia, ja), (lb, ib, jb) in pads:
        na, nb = b.new_internal_node(), b.new_internal_node()
        kinds = [KIND_RESISTOR] * 5
        nodes = [[na, nb, 0, 0],
                 [na, _pick(g[la], ia, ja), 0, 0], [na, _pick(g[la], ia + 1, ja), 0, 0],
                 [nb, _pick(g[lb], ib, jb), 0, 0], [nb, _pick(g[lb], ib + 1, jb), 0, 0]]
        b.add_mixed_network(kinds, nodes, [0.1, 1e-3, 1e-3, 1e-3, 1e-3])
    return b


CONFIGS = {
    "via_tht_4layer_standin": board_via_tht_standin,
    "2layer_1M": board_2layer,
    "4layer_regulator_4M": board_4layer_regulator,
    "8layer_10M": board_8layer,
}


# --------------------------------------------------------------------------------------------
# the same board as reference objects (build container only; small sizes)
# --------------------------------------------------------------------------------------------
def to_reference_objects(board: SynthBoard, problem_mod, mesh_mod, geometry_mod):
    """Build ``(prob, meshes, mesh_index_to_layer_index)`` out of the REFERENCE's own classes.

    ``problem_mod`` / ``mesh_mod`` are ``padne.problem`` / ``padne.mesh`` (for example from
    ``oracle.refload.load_reference()``); ``geometry_mod`` provides ``Point(x, y)`` and
    ``MultiPolygon(geoms)`` (shapely, or the loader's stub).  Vertex terminals become
    ``Connection``s placed exactly on the vertex so the reference's KD-tree lookup
    (``padne/solver.py:421-433``) resolves to that vertex; internal nodes become bare ``NodeID``s.
    Each network gets its own ``NodeID`` per distinct terminal, like the KiCad frontend does.
    """
    layers = [problem_mod.Layer(shape=geometry_mod.MultiPolygon(()), name=nm, conductance=c)
              for nm, c in zip(board.layer_names, board.layer_conductance)]
    meshes = []
    for xy, tri in board.meshes:
        pts = [mesh_mod.Point(float(x), float(y)) for x, y in xy]
        meshes.append(mesh_mod.Mesh.from_triangle_soup(pts, [tuple(int(i) for i in t) for t in tri]))
    v_off = board.v_off
    n_v = board.n_vertices
    all_xy = np.concatenate([m[0] for m in board.meshes])
    mesh_of = np.searchsorted(v_off, np.arange(n_v), side="right") - 1
    internal_ids: dict[int, object] = {}
    networks = []
    for net in board.networks:
        node_ids: dict[int, object] = {}
        connections = []

        def nid(t, node_ids=node_ids, connections=connections):
            t = int(t)
            if t in node_ids:
                return node_ids[t]
            if t < n_v:
                layer = layers[board.mesh_layer[int(mesh_of[t])]]
                conn = problem_mod.Connection(layer=layer,
                                              point=geometry_mod.Point(float(all_xy[t, 0]), float(all_xy[t, 1])))
                connections.append(conn)
                node_ids[t] = conn.node_id
            else:
                # an internal node is private to its network in the reference data model
                node_ids[t] = internal_ids.setdefault(t, problem_mod.NodeID())
            return node_ids[t]

        elements = []
        for k, nd, v, g in zip(net["kind"], net["node"], net["val"], net["gain"]):
            if k == KIND_RESISTOR:
                elements.append(problem_mod.Resistor(a=nid(nd[0]), b=nid(nd[1]), resistance=float(v)))
            elif k == KIND_CURRENT:
                elements.append(problem_mod.CurrentSource(f=nid(nd[0]), t=nid(nd[1]), current=float(v)))
            elif k == KIND_VOLTAGE:
                elements.append(problem_mod.VoltageSource(p=nid(nd[0]), n=nid(nd[1]), voltage=float(v)))
            elif k == KIND_REGULATOR:
                elements.append(problem_mod.VoltageRegulator(
                    v_p=nid(nd[0]), v_n=nid(nd[1]), s_f=nid(nd[2]), s_t=nid(nd[3]),
                    voltage=float(v), gain=float(g)))
            else:
                raise ValueError(k)
        networks.append(problem_mod.Network(connections=connections, elements=elements))
    prob = problem_mod.Problem(layers=layers, networks=networks, project_name=board.name)
    return prob, meshes, list(board.mesh_layer)
