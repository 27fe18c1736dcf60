"""ctypes front end of the CPU oracle (``oracle/padne_oracle.c``).  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import this module, and only as the checker or the timed CPU baseline.
``padne_b200/`` never imports it.

Parity status: PINNED (see the header of ``padne_oracle.c`` and ``tests/test_oracle.py``).

The linear solve of the reference is third-party code that is not vendored in ``/root/reference``:
``scipy.sparse.linalg.spsolve`` (SuperLU), declared as ``scipy>=1.15.0`` in the reference's
``pyproject.toml:46`` with no lock file; this image has scipy 1.18.1.  :func:`solve` calls it
exactly the way ``padne/solver.py:772-779`` does, so the oracle for the solve IS the reference's
own solver call on the identical matrix.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "padne_oracle.c")
_LIB = os.path.join(_HERE, "_build", "libpadne_oracle.so")
_lib = None

STATUS = {0: "ok", 1: "invalid argument", 2: "degenerate triangle (ZeroDivisionError)",
          7: "non-manifold mesh", 8: "out of memory"}


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__(f"oracle status {code}: {STATUS.get(code, '?')}")
        self.code = code


def build(force: bool = False) -> str:
    """Compile the C restatement with gcc (no FMA contraction).  Returns the .so path."""
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall",
                               "-o", _LIB, _SRC, "-lm"])
    return _LIB


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(build())
        P = ctypes.c_void_p
        lib.oracle_assemble.argtypes = [ctypes.c_int32, P, P, P, P, P, ctypes.c_int64, ctypes.c_int64,
                                        P, P, P, P, ctypes.POINTER(P), ctypes.POINTER(ctypes.c_int64)]
        lib.oracle_system_csr.argtypes = [P, P, P, P]
        lib.oracle_system_free.argtypes = [P]
        lib.oracle_system_free.restype = None
        lib.oracle_corner_weights.argtypes = [ctypes.c_int64, P, ctypes.c_int64, P, P]
        lib.oracle_power_density.argtypes = [ctypes.c_int64, P, P, ctypes.c_int64, P, ctypes.c_double, P]
        _lib = lib
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def assemble(fs, stamps):
    """``assemble_system`` (``solver.py:783-812``) -> ``(indptr, indices, data)`` of ``L.tocsr()``.

    ``fs`` is a ``padne_b200.flat.FlatSystem``; ``stamps`` is ``(s_row, s_col, s_val, s_op)`` from
    ``padne_b200.stamps.build_stamps``.
    """
    lib = _load()
    s_row, s_col, s_val, s_op = [np.ascontiguousarray(a) for a in stamps[:4]]
    h = ctypes.c_void_p()
    nnz = ctypes.c_int64()
    rc = lib.oracle_assemble(fs.n_mesh, _p(fs.v_off), _p(fs.xy), _p(fs.t_off), _p(fs.tri),
                             _p(fs.conductance), fs.N, s_row.shape[0], _p(s_row), _p(s_col),
                             _p(s_val), _p(s_op), ctypes.byref(h), ctypes.byref(nnz))
    if rc:
        raise OracleError(rc)
    try:
        indptr = np.empty(fs.N + 1, np.int32)
        indices = np.empty(nnz.value, np.int32)
        data = np.empty(nnz.value, np.float64)
        lib.oracle_system_csr(h, _p(indptr), _p(indices), _p(data))
    finally:
        lib.oracle_system_free(h)
    return indptr, indices, data


def corner_weights(xy, tri):
    """Per-triangle, per-corner ``abs(dot/cross)/2`` (``mesh.py:136-138``); ``tri`` indexes ``xy``."""
    lib = _load()
    xy = np.ascontiguousarray(xy, np.float64)
    tri = np.ascontiguousarray(tri, np.int32)
    w = np.empty((tri.shape[0], 3), np.float64)
    rc = lib.oracle_corner_weights(xy.shape[0], _p(xy), tri.shape[0], _p(tri), _p(w))
    if rc:
        raise OracleError(rc)
    return w


def power_density(xy, values, tri, conductivity):
    """``compute_power_density`` (``solver.py:728-745``) per face."""
    lib = _load()
    xy = np.ascontiguousarray(xy, np.float64)
    values = np.ascontiguousarray(values, np.float64)
    tri = np.ascontiguousarray(tri, np.int32)
    p = np.empty(tri.shape[0], np.float64)
    rc = lib.oracle_power_density(xy.shape[0], _p(xy), _p(values), tri.shape[0], _p(tri),
                                  float(conductivity), _p(p))
    if rc:
        raise OracleError(rc)
    return p


def solve(indptr, indices, data, r):
    """``solve_system`` (``solver.py:767-780``): ``(v, residual_norm, ground_node_current)``."""
    import scipy.sparse
    import scipy.sparse.linalg
    n = len(indptr) - 1
    L_csc = scipy.sparse.csr_matrix((data, indices, indptr), shape=(n, n)).tocsc()
    v = scipy.sparse.linalg.spsolve(L_csc, r)
    residual_norm = float(np.linalg.norm(L_csc @ v - r))
    return v, residual_norm, float(v[-1])


def assemble_and_solve(fs, stamps_and_rhs):
    """The whole reference hot path on one flat system (used as the timed CPU baseline 'port')."""
    indptr, indices, data = assemble(fs, stamps_and_rhs[:4])
    v, res, ig = solve(indptr, indices, data, stamps_and_rhs[4])
    return (indptr, indices, data), v, res, ig


def nearest_vertex(v_off, mesh_layer, xy, q_xy, q_layer):
    """Row F3: what ``NodeIndexer.create`` gets from its per-layer KD-trees (``padne/solver.py:356-396``
    build, ``:421-433`` query), restated as a brute-force search in numpy: per query the vertex with the
    smallest ``dx*dx + dy*dy`` among all meshes of the query's layer, lowest global index on an exact
    tie, -1 when the layer has no vertex.  ``tests/test_oracle.py`` pins it against
    ``scipy.spatial.KDTree(..., leafsize=32).query`` -- the reference's own call -- on tie-free inputs.
    Returns (index int64 [n_q], squared distance [n_q])."""
    v_off = np.asarray(v_off, np.int64)
    mesh_layer = np.asarray(mesh_layer, np.int32)
    xy = np.asarray(xy, np.float64).reshape(-1, 2)
    q_xy = np.asarray(q_xy, np.float64).reshape(-1, 2)
    q_layer = np.asarray(q_layer, np.int32)
    v_layer = np.repeat(mesh_layer, np.diff(v_off))
    out = np.full(q_xy.shape[0], -1, np.int64)
    d2 = np.full(q_xy.shape[0], np.inf)
    for layer in np.unique(q_layer):
        idx = np.flatnonzero(v_layer == layer)
        if idx.size == 0:
            continue
        p = xy[idx]
        for q in np.flatnonzero(q_layer == layer):
            dx = p[:, 0] - q_xy[q, 0]
            dy = p[:, 1] - q_xy[q, 1]
            d = dx * dx + dy * dy
            k = int(np.argmin(d))            # first minimum == lowest global index (idx is ascending)
            out[q] = idx[k]
            d2[q] = d[k]
    return out, d2


# -- row F4: the reference's .vtu element tree, restated over arrays (pure Python, small cases) ------------
def vtu_piece_tree(xy, tri, values):
    """``[tag, [[attr, value], ...], text, [children]]`` of one ``<Piece>``: ``create_piece``
    ``padne/paraview.py:209-231`` with its children ``create_point_data`` (``:109-126``), ``create_points``
    (``:129-146``, y negated, z = 0.0) and ``create_cells`` (``:184-206``); every payload is
    ``" ".join(str(value) ...)`` over Python floats / ints like ``create_data_array`` (``:62-93``)."""
    xy = [(float(x), float(y)) for x, y in np.asarray(xy, np.float64).reshape(-1, 2)]
    tri = [tuple(int(i) for i in t) for t in np.asarray(tri).reshape(-1, 3)]
    values = [float(x) for x in np.asarray(values, np.float64).reshape(-1)]

    def data_array(data_type, items, name=None, n_comp=None):
        attrs = [["type", data_type], ["format", "ascii"]]
        if name is not None:
            attrs.append(["Name", name])
        if n_comp is not None:
            attrs.append(["NumberOfComponents", str(n_comp)])
        return ["DataArray", attrs, " ".join(str(v) for v in items), []]

    coords = []
    for x, y in xy:
        coords.extend([x, -y, 0.0])
    conn = [i for t in tri for i in t]
    return ["Piece", [["NumberOfPoints", str(len(xy))], ["NumberOfCells", str(len(tri))]], None, [
        ["PointData", [["Scalars", "voltage"]], None, [data_array("Float64", values, name="voltage")]],
        ["Points", [], None, [data_array("Float64", coords, n_comp=3)]],
        ["Cells", [], None, [data_array("Int32", conn, name="connectivity"),
                             data_array("Int32", [3 * (i + 1) for i in range(len(tri))], name="offsets"),
                             data_array("UInt8", [5] * len(tri), name="types")]]]]


def vtu_file_tree(pieces):
    """One layer document (``create_vtk_root`` ``padne/paraview.py:96-106`` + ``export_solution`` ``:270-281``)."""
    return ["VTKFile", [["type", "UnstructuredGrid"], ["version", "0.1"], ["byte_order", "LittleEndian"]], None,
            [["UnstructuredGrid", [], None, [vtu_piece_tree(*p) for p in pieces]]]]
