"""Flat (array) description of one padne FEM system -- what crosses the C ABI.

The reference walks Python object graphs (``mesh.Mesh`` half-edges, ``problem.Network``
element lists).  The kernels need arrays.  ``FlatSystem`` is the array form of exactly the
inputs ``padne/solver.py:783-812`` (``assemble_system``) consumes:

* meshes laid out back to back like ``VertexIndexer`` does (``solver.py:216-229``):
  mesh ``m`` owns global vertex indices ``[v_off[m], v_off[m+1])``;
* per-mesh conductance = the owning layer's conductance (``solver.py:796-799``);
* the lumped elements of ``filtered_networks`` in network/element order with their
  terminals already resolved to global indices by the reference's own ``NodeIndexer``
  (``solver.py:350-466``) -- never recomputed here, because internal-node numbering depends
  on ``set`` iteration over identity-hashed ``NodeID`` objects (``problem.py:86-96``);
* the ground vertex chosen by ``find_best_ground_node_index`` (``solver.py:671-686``).

Unknown order (``solver.py:757-760``):
``[mesh vertices | internal nodes | one current per VoltageSource/VoltageRegulator | ground current]``.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

# element kinds (``problem.py:104-174``)
KIND_RESISTOR = 0
KIND_CURRENT = 1
KIND_VOLTAGE = 2
KIND_REGULATOR = 3

# stamp operations: LIL ``+=`` versus ``=`` (``solver.py:480-484`` vs ``solver.py:503-511``)
OP_ADD = 0
OP_SET = 1


@dataclass
class FlatSystem:
    # geometry -----------------------------------------------------------------------
    v_off: np.ndarray            # int64 [n_mesh+1]
    t_off: np.ndarray            # int64 [n_mesh+1]
    xy: np.ndarray               # float64 [n_vertices, 2]
    tri: np.ndarray              # int32 [n_triangles, 3], MESH-LOCAL vertex indices
    conductance: np.ndarray      # float64 [n_mesh]
    # unknown bookkeeping ------------------------------------------------------------
    n_internal: int = 0          # NodeIndexer.internal_node_count
    n_extra: int = 0             # len(NodeIndexer.extra_source_to_global_index)
    i_gnd: int = 0               # find_best_ground_node_index
    # elements in stamping order -----------------------------------------------------
    # Resistor:  (a, b, -, -)   val=resistance
    # Current:   (f, t, -, -)   val=current
    # Voltage:   (p, n, -, -)   val=voltage            xidx = i_v
    # Regulator: (v_p, v_n, s_f, s_t) val=voltage gain xidx = i_v
    e_kind: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    e_node: np.ndarray = field(default_factory=lambda: np.zeros((0, 4), np.int32))
    e_val: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float64))
    e_gain: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float64))
    e_xidx: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    name: str = ""

    def __post_init__(self):
        self.v_off = np.ascontiguousarray(self.v_off, dtype=np.int64)
        self.t_off = np.ascontiguousarray(self.t_off, dtype=np.int64)
        self.xy = np.ascontiguousarray(self.xy, dtype=np.float64).reshape(-1, 2)
        self.tri = np.ascontiguousarray(self.tri, dtype=np.int32).reshape(-1, 3)
        self.conductance = np.ascontiguousarray(self.conductance, dtype=np.float64)
        self.e_kind = np.ascontiguousarray(self.e_kind, dtype=np.int32)
        self.e_node = np.ascontiguousarray(self.e_node, dtype=np.int32).reshape(-1, 4)
        self.e_val = np.ascontiguousarray(self.e_val, dtype=np.float64)
        self.e_gain = np.ascontiguousarray(self.e_gain, dtype=np.float64)
        self.e_xidx = np.ascontiguousarray(self.e_xidx, dtype=np.int32)
        if self.v_off.shape != self.t_off.shape or self.v_off.shape[0] != self.conductance.shape[0] + 1:
            raise ValueError("v_off / t_off / conductance length mismatch")
        if int(self.v_off[-1]) != self.xy.shape[0] or int(self.t_off[-1]) != self.tri.shape[0]:
            raise ValueError("offset tables do not cover xy / tri")

    # sizes (solver.py:757-760) ------------------------------------------------------
    @property
    def n_mesh(self) -> int:
        return int(self.conductance.shape[0])

    @property
    def n_vertices(self) -> int:
        return int(self.xy.shape[0])

    @property
    def n_triangles(self) -> int:
        return int(self.tri.shape[0])

    @property
    def n_nodes(self) -> int:
        """Voltage unknowns: mesh vertices followed by internal network nodes."""
        return self.n_vertices + int(self.n_internal)

    @property
    def N(self) -> int:
        return self.n_nodes + int(self.n_extra) + 1

    @property
    def n_elements(self) -> int:
        return int(self.e_kind.shape[0])

    @property
    def has_regulator(self) -> bool:
        return bool(np.any(self.e_kind == KIND_REGULATOR))

    # (de)serialisation for the golden fixtures -----------------------------------------
    _ARRAYS = ("v_off", "t_off", "xy", "tri", "conductance",
               "e_kind", "e_node", "e_val", "e_gain", "e_xidx")

    def to_npz_dict(self, prefix: str = "sys_") -> dict:
        d = {prefix + k: getattr(self, k) for k in self._ARRAYS}
        d[prefix + "scalars"] = np.array([self.n_internal, self.n_extra, self.i_gnd], np.int64)
        return d

    @classmethod
    def from_npz_dict(cls, d, prefix: str = "sys_", name: str = "") -> "FlatSystem":
        sc = d[prefix + "scalars"]
        return cls(**{k: d[prefix + k] for k in cls._ARRAYS},
                   n_internal=int(sc[0]), n_extra=int(sc[1]), i_gnd=int(sc[2]), name=name)
