"""Host-side mirror of the reference's solver seams (``padne/solver.py``) over the C ABI.

Same names, argument meaning and error behaviour as the reference so that callers and tests read
the same:

==============================  ==================================  ==========================
this module                     replaces (atx/padne)                 runs on
==============================  ==================================  ==========================
``laplace_operator``            ``solver.py:171-213``                kernels (1)+(2)
``assemble_system``             ``solver.py:783-812``                kernels (1)+(2)
``solve_system``                ``solver.py:767-780``                kernels (3)+(4)
``stamp_network_into_system``   ``solver.py:469-541``                host (op list, same writes)
``compute_power_density``       ``solver.py:728-745``                power-density kernel
``produce_layer_solutions``     ``solver.py:578-615``                power-density kernel + copies
``solve``                       ``solver.py:815-902``                all of the above
==============================  ==================================  ==========================

The pre-passes that stay the reference's own code (connectivity, meshing, ``VertexIndexer``,
``filter_dead_networks``, ``NodeIndexer`` -- SURVEY.md section 7) are called through the
``padne.solver`` module of the installed reference; nothing here recomputes them.

There is no CPU fallback: every function that computes goes through :mod:`padne_b200.capi` and
raises if the CUDA library or a GPU is missing.
"""
from __future__ import annotations

import logging
import warnings
from dataclasses import dataclass
from typing import Optional

import numpy as np
import scipy.sparse

from . import capi, halfedge
from .flat import FlatSystem, OP_SET
from .flatten import flatten_elements, flatten_system, mesh_to_arrays
from .stamps import build_stamps

log = logging.getLogger(__name__)

DTYPE = np.float64       # padne/solver.py:21


@dataclass
class SolverInfo:
    """Mirror of ``padne/solver.py:33-37`` (used when the reference module is not importable)."""
    ground_node_current: float
    residual_norm: float


class SolverWarning(Warning):
    """Mirror of the reference's ``SolverWarning`` (``padne/solver.py:28-30``)."""


# solver knobs: optional, never positional (the reference's solve() has no options)
@dataclass
class SolverOptions:
    rtol: float = 1e-10      # required relative residual of the reduced system (north star)
    atol: float = 0.0        # absolute bar on ||L v - r||_2; 0 -> the reference's own 1e-9
                             # (tests/test_solver.py:2083-2089), < 0 -> off
    max_iter: int = 0
    method: int = capi.METHOD_AUTO
    precond: int = capi.PRECOND_AUTO
    device: int = 0


OPTIONS = SolverOptions()


def _reference_solver_module():
    """The reference's own ``padne.solver`` if it is importable (drop-in deployment), else None."""
    try:
        import padne.solver as ref  # noqa: PLC0415
        return ref
    except Exception:
        return None


class ResidentSystem(scipy.sparse.csr_matrix):
    """``L`` as returned by :func:`assemble_system`: a ``csr_matrix`` (so ``L.shape``, ``L.nnz``,
    ``L.tocsc()`` keep working, ``benchmarks/benchmarks.py:652-654``) that remembers that the same
    matrix is still resident on the GPU, which lets :func:`solve_system` skip the upload."""
    _pdn_ctx = None
    _pdn_token = None
    _pdn_n_nodes = -1


_resident_token = [0]


# ------------------------------------------------------------------------------------------------
# kernel (1)+(2) seams
# ------------------------------------------------------------------------------------------------
def laplace_operator(mesh, ctx: Optional[capi.Context] = None) -> scipy.sparse.coo_matrix:
    """``laplace_operator(mesh)`` (``padne/solver.py:171-213``): the per-mesh cotangent operator.

    Negative semidefinite, entries whose summed weight is exactly 0 are absent
    (``solver.py:187-190``).  Returned as a ``coo_matrix`` like the reference (row-major sorted
    entry order; the reference emits off-diagonals in orbit order then diagonals -- the matrices
    are equal, bit for bit, entry by entry)."""
    ctx = ctx or capi.default_context(OPTIONS.device)
    xy, tri = mesh_to_arrays(mesh)
    n = xy.shape[0]
    if n == 0:
        return scipy.sparse.coo_matrix((0, 0), dtype=DTYPE)
    z32 = np.zeros(0, np.int32)
    ctx.assemble(np.array([0, n], np.int64), xy, np.array([0, tri.shape[0]], np.int64), tri,
                 np.array([1.0]), n, z32, z32, np.zeros(0), z32)
    indptr, indices, data = ctx.get_csr()
    return scipy.sparse.csr_matrix((data, indices, indptr), shape=(n, n)).tocoo()


def assemble_flat(fs: FlatSystem, ctx: Optional[capi.Context] = None):
    """Assemble a :class:`FlatSystem`; returns ``(L, r)`` with ``L`` resident on the GPU."""
    ctx = ctx or capi.default_context(OPTIONS.device)
    s_row, s_col, s_val, s_op, rhs = build_stamps(fs)
    ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, s_row, s_col, s_val, s_op)
    indptr, indices, data = ctx.get_csr()
    L = ResidentSystem((data, indices, indptr), shape=(fs.N, fs.N))
    _resident_token[0] += 1
    L._pdn_ctx, L._pdn_token, L._pdn_n_nodes = ctx, _resident_token[0], fs.n_nodes
    ctx._resident_token = _resident_token[0]
    return L, rhs


def assemble_system(prob, meshes, mesh_index_to_layer_index, vindex, filtered_networks, node_indexer,
                    ctx: Optional[capi.Context] = None):
    """``assemble_system`` (``padne/solver.py:783-812``) -> ``(L, r)``.

    ``L`` is a ``scipy.sparse`` matrix equal to the reference's ``L.tocsr()`` (same sorted int32
    pattern, same values); ``r`` is the right-hand side."""
    fs = flatten_system(prob, meshes, mesh_index_to_layer_index, vindex, filtered_networks, node_indexer,
                        name=getattr(prob, "project_name", ""))
    return assemble_flat(fs, ctx)


def stamp_network_into_system(network, node_indexer, L, r) -> None:
    """``stamp_network_into_system`` (``padne/solver.py:469-541``): mutates ``L`` (any scipy matrix
    supporting item assignment, e.g. ``lil_matrix``) and ``r`` in place with the same ordered
    writes as the reference (``+=`` for resistor / gain stamps, ``=`` for source stamps)."""
    kind, node, val, gain, xidx = flatten_elements([network], node_indexer)
    N = L.shape[0]
    fs = FlatSystem(v_off=np.zeros(1, np.int64), t_off=np.zeros(1, np.int64), xy=np.zeros((0, 2)),
                    tri=np.zeros((0, 3), np.int32), conductance=np.zeros(0),
                    n_internal=N - 1, n_extra=0, i_gnd=0,
                    e_kind=kind, e_node=node, e_val=val, e_gain=gain, e_xidx=xidx)
    s_row, s_col, s_val, s_op, rhs = build_stamps(fs, with_ground=False)
    for i, j, x, op in zip(s_row.tolist(), s_col.tolist(), s_val.tolist(), s_op.tolist()):
        if op == OP_SET:
            L[i, j] = x
        else:
            L[i, j] += x
    _apply_rhs(fs, r)


def _apply_rhs(fs: FlatSystem, r: np.ndarray) -> None:
    """RHS writes of one network onto an existing ``r`` (``solver.py:489-492, 505, 530``)."""
    from .flat import KIND_CURRENT, KIND_REGULATOR, KIND_VOLTAGE
    for k, nd, v, xi in zip(fs.e_kind.tolist(), fs.e_node.tolist(), fs.e_val.tolist(), fs.e_xidx.tolist()):
        if k == KIND_CURRENT:
            r[nd[0]] += v
            r[nd[1]] -= v
        elif k == KIND_VOLTAGE:
            r[xi] = v
        elif k == KIND_REGULATOR:
            r[xi] += v


# ------------------------------------------------------------------------------------------------
# kernel (3)+(4) seam
# ------------------------------------------------------------------------------------------------
def solve_system(L, r, ctx: Optional[capi.Context] = None, n_nodes: Optional[int] = None, **knobs):
    """``solve_system(L, r)`` (``padne/solver.py:767-780``) -> ``(v, SolverInfo)``.

    ``residual_norm`` is ``||L v - r||_2`` of the full system and ``ground_node_current`` is
    ``v[-1]``, as in the reference.  A structurally singular constraint set behaves like the
    reference's SuperLU path: a ``MatrixRankWarning`` and an all-NaN ``v``."""
    v, info = _solve_system_info(L, r, ctx, n_nodes, **knobs)
    ref = _reference_solver_module()
    cls = ref.SolverInfo if ref is not None else SolverInfo
    return v, cls(ground_node_current=float(info.ground_node_current) if info else float("nan"),
                  residual_norm=float(info.residual_norm) if info else float("nan"))


last_solve_info: Optional[capi.SolveInfo] = None


def _solve_system_info(L, r, ctx=None, n_nodes=None, **knobs):
    global last_solve_info
    resident = (isinstance(L, ResidentSystem) and L._pdn_ctx is not None and (ctx is None or ctx is L._pdn_ctx)
                and getattr(L._pdn_ctx, "_resident_token", None) == L._pdn_token)
    if resident:
        ctx = L._pdn_ctx
        if n_nodes is None:
            n_nodes = L._pdn_n_nodes
    else:
        ctx = ctx or capi.default_context(OPTIONS.device)
        csr = scipy.sparse.csr_matrix(L)
        csr.sum_duplicates()
        csr.sort_indices()
        csr.eliminate_zeros()
        ctx.set_csr(csr.indptr.astype(np.int32), csr.indices.astype(np.int32), csr.data.astype(np.float64))
        ctx._resident_token = None
    opts = dict(rtol=OPTIONS.rtol, atol=OPTIONS.atol, max_iter=OPTIONS.max_iter, method=OPTIONS.method,
                precond=OPTIONS.precond)
    opts.update(knobs)
    try:
        v, info = ctx.solve(np.asarray(r, dtype=np.float64), n_nodes=-1 if n_nodes is None else int(n_nodes), **opts)
    except capi.SingularSystemError as e:
        from scipy.sparse.linalg import MatrixRankWarning
        warnings.warn(f"Matrix is exactly singular ({e})", MatrixRankWarning)
        last_solve_info = None
        return np.full(L.shape[0], np.nan), None
    last_solve_info = info
    log.info("Krylov solve: %s, %d iterations, relative residual %.3e, %.1f ms", info.method,
             info.iterations, info.rel_residual, info.t_total_ms)
    # the reference is a direct solve: say so when the iteration stopped at the floor of the arithmetic above
    # the relative tolerance that was asked for (the C ABI has already refused anything worse than 2x that floor)
    rtol = opts.get("rtol") or 1e-10
    if info.stagnated and info.rel_residual > rtol * 1.000001:
        warnings.warn(f"Krylov solve stagnated at relative residual {info.rel_residual:.3e} (asked for {rtol:.1e}): "
                      f"fp64 floor {info.floor_abs:.2e} on a right-hand side of norm {info.rhs_norm:.2e}", SolverWarning)
    return v, info


# ------------------------------------------------------------------------------------------------
# row F1: power density / layer solutions
# ------------------------------------------------------------------------------------------------
def compute_power_density(voltage, conductivity: float, ctx: Optional[capi.Context] = None):
    """``compute_power_density`` (``padne/solver.py:728-745``): a ``TwoForm`` on ``voltage.mesh``."""
    ctx = ctx or capi.default_context(OPTIONS.device)
    msh = voltage.mesh
    xy, tri = mesh_to_arrays(msh)
    two_form_cls = _form_classes(voltage)[1]
    pd = two_form_cls(msh)
    if tri.shape[0]:
        pd.values[:] = ctx.power_density(xy, np.asarray(voltage.values, np.float64), tri, float(conductivity))
    return pd


def _form_classes(zero_form):
    import importlib
    mod = importlib.import_module(type(zero_form).__module__)
    return mod.ZeroForm, mod.TwoForm


def produce_layer_solutions(layers, vindex, meshes, mesh_index_to_layer_index, v,
                            disconnected_meshes_by_layer, ref=None, ctx: Optional[capi.Context] = None):
    """``produce_layer_solutions`` (``padne/solver.py:578-615``) built from the reference's own
    ``LayerSolution`` / ``ZeroForm`` / ``TwoForm`` classes, so ``ui``, ``paraview`` and pickle are
    unaffected.  One ``LayerSolution`` per layer, including layers without a mesh."""
    ref = ref or _reference_solver_module()
    if ref is None:
        raise RuntimeError("produce_layer_solutions needs the reference's padne.solver / padne.mesh classes")
    ctx = ctx or capi.default_context(OPTIONS.device)
    mesh_mod = ref.mesh
    offsets = np.concatenate([[0], np.cumsum([len(m.vertices) for m in meshes])]).astype(np.int64)
    out = []
    for layer_i, layer in enumerate(layers):
        l_meshes, l_values, l_pd = [], [], []
        for mesh_i, msh in enumerate(meshes):
            if mesh_index_to_layer_index[mesh_i] != layer_i:
                continue
            zf = mesh_mod.ZeroForm(msh)
            zf.values[:] = v[offsets[mesh_i]:offsets[mesh_i + 1]]     # global indices are back to back
            l_values.append(zf)
            l_meshes.append(msh)
            l_pd.append(compute_power_density(zf, layer.conductance, ctx))
        out.append(ref.LayerSolution(meshes=l_meshes, potentials=l_values, power_densities=l_pd,
                                     disconnected_meshes=disconnected_meshes_by_layer[layer_i]))
    return out


# ------------------------------------------------------------------------------------------------
# row F3: NodeIndexer with the nearest-vertex lookup on the GPU
# ------------------------------------------------------------------------------------------------
class _PrecomputedTree:
    """Stands in for one per-layer ``scipy.spatial.KDTree`` inside the reference's ``NodeIndexer.create``:
    ``query((x, y), k=1)`` answers from the batch that ``pdn_nearest_vertex`` already computed."""

    def __init__(self):
        self.answers = {}

    def query(self, point, k=1):
        return 0.0, self.answers[(float(point[0]), float(point[1]))]


class _IdentityIndex:
    """``layer_global_index_and_vertex[layer][i][0]`` of the reference, for trees that already answer with the
    global vertex index."""

    def __getitem__(self, i):
        return (int(i), None)


def node_indexer_create(ref, prob, meshes, mesh_index_to_layer_index, vindex, filtered_networks,
                        ctx: Optional[capi.Context] = None, lookup=None):
    """``NodeIndexer.create`` (``padne/solver.py:399-466``) with its KD-trees replaced by ONE call of
    ``pdn_nearest_vertex`` for all connections.

    The reference's own ``create`` still runs -- the duplicate check and the numbering of the internal
    nodes (which follows ``set`` iteration over identity-hashed ``NodeID``s) stay its code; only
    ``_construct_kdtrees`` (``:356-396``) is swapped, for the duration of the call, for objects that
    answer ``kdtree.query`` (``:424``) from the precomputed batch.  ``lookup(v_off, mesh_layer, xy, q_xy,
    q_layer) -> global indices`` defaults to the GPU; tests inject the oracle."""
    if lookup is None:
        lookup = (ctx or capi.default_context(OPTIONS.device)).nearest_vertex
    conns = [c for net in filtered_networks for c in net.connections]
    v_off = np.concatenate([[0], np.cumsum([len(m.vertices) for m in meshes])]).astype(np.int64)
    xy = (np.concatenate([mesh_to_arrays(m)[0] for m in meshes]) if len(meshes)
          else np.zeros((0, 2)))                             # one walk per mesh, shared with the assembly (row F2)
    q_xy = np.array([(c.point.x, c.point.y) for c in conns], np.float64).reshape(-1, 2)
    q_layer = np.array([prob.layers.index(c.layer) for c in conns], np.int32)
    idx = lookup(v_off, np.asarray(mesh_index_to_layer_index, np.int32), xy, q_xy, q_layer) if len(conns) else []
    trees, index_lists = {}, {}
    for layer_i in set(int(l) for l in mesh_index_to_layer_index):
        if any(len(m.vertices) for m, l in zip(meshes, mesh_index_to_layer_index) if l == layer_i):
            trees[layer_i] = _PrecomputedTree()           # an empty layer gets none, like the reference (:381-386)
            index_lists[layer_i] = _IdentityIndex()
    for c, layer_i, g in zip(conns, q_layer, idx):
        if int(g) >= 0:
            trees[int(layer_i)].answers[(float(c.point.x), float(c.point.y))] = int(g)
    original = ref.NodeIndexer.__dict__["_construct_kdtrees"]
    ref.NodeIndexer._construct_kdtrees = classmethod(lambda cls, *a, **k: (trees, index_lists))
    try:
        return ref.NodeIndexer.create(prob, meshes, mesh_index_to_layer_index, vindex, filtered_networks)
    finally:
        ref.NodeIndexer._construct_kdtrees = original


# ------------------------------------------------------------------------------------------------
# the public entry point
# ------------------------------------------------------------------------------------------------
def solve_meshed(prob, meshes, mesh_index_to_layer_index, disconnected_meshes_by_layer=None,
                 filtered_networks=None, ref=None, ctx: Optional[capi.Context] = None):
    """Steps 4-10 of ``solve`` (``padne/solver.py:860-902``) for already meshed layers."""
    ref = ref or _reference_solver_module()
    if ref is None:
        raise RuntimeError("padne_b200.solver.solve needs the reference's padne package for the host pre-passes")
    if disconnected_meshes_by_layer is None:
        disconnected_meshes_by_layer = [[] for _ in prob.layers]
    log.info("Indexing vertices and connections")
    vindex = ref.VertexIndexer.create(meshes)
    if filtered_networks is None:
        filtered_networks = list(prob.networks)
    log.info("Constructing node index for networks")
    node_indexer = node_indexer_create(ref, prob, meshes, mesh_index_to_layer_index, vindex, filtered_networks, ctx)
    log.info("Assembling the global system")
    L, r = assemble_system(prob, meshes, mesh_index_to_layer_index, vindex, filtered_networks, node_indexer, ctx)
    log.info("Solving the system of equations")
    v, solver_info = solve_system(L, r, ctx)
    if not np.isclose(solver_info.ground_node_current, 0):
        warnings.warn(
            f"Ground node current is not zero ({solver_info.ground_node_current} A), this may indicate an issue "
            "with the problem being solved. Check for unterminated current loops or floating connected "
            "components. This may be harmless if the current is small, but it may indicate an "
            "ill-conditioned system.", ref.SolverWarning)
    log.info("Producing the solution object")
    layer_solutions = produce_layer_solutions(prob.layers, vindex, meshes, mesh_index_to_layer_index, v,
                                              disconnected_meshes_by_layer, ref, ctx)
    return ref.Solution(problem=prob, layer_solutions=layer_solutions, solver_info=solver_info)


def solve(prob, mesher_config=None):
    """``solve(prob, mesher_config=None) -> Solution`` (``padne/solver.py:815``), same signature.

    Connectivity, meshing and dead-network filtering are the reference's own functions
    (``solver.py:245, 263, 321, 654``); assembly, solve and power density run on the GPU."""
    ref = _reference_solver_module()
    if ref is None:
        raise RuntimeError("padne_b200.solver.solve needs the reference's padne package for the host pre-passes")
    mesher = ref.mesh.Mesher(mesher_config)
    log.info("Constructing connectivity graph and finding connected layers")
    strtrees, _, connected = ref.compute_connectivity(prob)
    log.info("Meshing the connected components")
    with halfedge.fast_triangle_soup(ref.mesh):              # row F2: the mesher's soup -> Mesh objects + flat arrays
        meshes, m2l = ref.generate_meshes_for_problem(prob, mesher, connected, strtrees)
    log.info("Meshing the disconnected components")
    with halfedge.fast_triangle_soup(ref.mesh):
        disconnected = ref.generate_disconnected_meshes(prob, connected)
    log.info("Processing lumped element networks")
    filtered = ref.filter_dead_networks(prob, strtrees, connected)
    log.info(f"Filtered networks: {len(filtered)}/{len(prob.networks)}")
    return solve_meshed(prob, meshes, m2l, disconnected, filtered, ref)
