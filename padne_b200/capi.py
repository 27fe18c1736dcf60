"""ctypes binding of ``libpadne_b200.so`` (``include/padne_b200.h``) -- the only door into the kernels.

There is NO CPU fallback: if the library is missing, or no CUDA device answers, every compute
entry point raises.  Status codes are mapped to the exception types the reference raises
(error conventions).
"""
from __future__ import annotations

import ctypes
import os
import warnings
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libpadne_b200.so")

PDN_OK, PDN_EINVAL, PDN_EDEGENERATE, PDN_ESINGULAR, PDN_ENOTCONVERGED = 0, 1, 2, 3, 4
PDN_ECUDA, PDN_EUNSUPPORTED, PDN_ENONMANIFOLD, PDN_ENOMEM, PDN_ENOSYSTEM = 5, 6, 7, 8, 9
METHOD_AUTO, METHOD_PCG, METHOD_BICGSTAB = 0, 1, 2
PRECOND_AUTO, PRECOND_NONE, PRECOND_TWOLEVEL, PRECOND_JACOBI = 0, 1, 3, 4
PRECOND_NAMES = {PRECOND_NONE: "none", PRECOND_TWOLEVEL: "two-level", PRECOND_JACOBI: "jacobi"}
METHOD_NAMES = {METHOD_PCG: "pcg", METHOD_BICGSTAB: "bicgstab"}

# every symbol include/padne_b200.h declares (tests check that the library exports them all)
SYMBOLS = ("pdn_create", "pdn_destroy", "pdn_strerror", "pdn_last_error", "pdn_launch_count",
           "pdn_stream", "pdn_element_weights", "pdn_assemble", "pdn_assemble_part", "pdn_get_csr", "pdn_get_shape",
           "pdn_set_csr", "pdn_solve", "pdn_spmv", "pdn_bench_spmv", "pdn_residual_noalloc", "pdn_power_density",
           "pdn_comm_unique_id", "pdn_comm_init", "pdn_comm_destroy", "pdn_comm_info", "pdn_nearest_vertex",
           "pdn_format_f64", "pdn_vtu_piece_bound", "pdn_vtu_piece")


class SingularSystemError(RuntimeError):
    """The constraint structure is singular (the reference: SuperLU warning + all-NaN ``v``)."""


class NotConvergedError(RuntimeError):
    """The Krylov solver hit ``max_iter`` -- never silent."""


class PdnOpts(ctypes.Structure):
    _fields_ = [("rtol", ctypes.c_double), ("atol", ctypes.c_double), ("max_iter", ctypes.c_int32),
                ("method", ctypes.c_int32), ("precond", ctypes.c_int32), ("check_every", ctypes.c_int32),
                ("n_nodes", ctypes.c_int64), ("n_coarse", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class PdnInfo(ctypes.Structure):
    _fields_ = [("method", ctypes.c_int32), ("iterations", ctypes.c_int32), ("restarts", ctypes.c_int32),
                ("n_constraints", ctypes.c_int32), ("rel_residual", ctypes.c_double),
                ("residual_norm", ctypes.c_double), ("ground_node_current", ctypes.c_double),
                ("t_setup_ms", ctypes.c_double), ("t_krylov_ms", ctypes.c_double),
                ("t_total_ms", ctypes.c_double), ("spmv_ms", ctypes.c_double),
                ("precond", ctypes.c_int32), ("n_coarse", ctypes.c_int32), ("t_coarse_ms", ctypes.c_double),
                ("n_coarse_local", ctypes.c_int32), ("local_level", ctypes.c_int32), ("stagnated", ctypes.c_int32),
                ("iteration_kind", ctypes.c_int32), ("abs_target", ctypes.c_double), ("floor_abs", ctypes.c_double),
                ("rhs_norm", ctypes.c_double), ("spmv_kind", ctypes.c_int32), ("spmv_halo", ctypes.c_int32),
                ("spmv_window_coverage", ctypes.c_double)]


@dataclass
class SolveInfo:
    method: str
    iterations: int
    restarts: int
    n_constraints: int
    rel_residual: float
    residual_norm: float
    ground_node_current: float
    t_setup_ms: float
    t_krylov_ms: float
    t_total_ms: float
    precond: str = "jacobi"
    n_coarse: int = 0
    t_coarse_ms: float = 0.0
    n_coarse_local: int = 0
    local_level: int = 0
    stagnated: int = 0
    iteration_kind: int = 1
    abs_target: float = 0.0
    floor_abs: float = 0.0
    rhs_norm: float = 0.0
    spmv_kind: int = 0
    spmv_halo: int = 0
    spmv_window_coverage: float = 0.0


_lib = None


def load_library(path: str | None = None):
    """Load the shared library (no GPU needed for loading)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("PADNE_B200_LIB") or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(f"{p} is missing: run `python -m padne_b200.build` (there is no CPU fallback)")
    lib = ctypes.CDLL(p)
    P, I32, I64, D = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_double
    lib.pdn_create.argtypes = [ctypes.POINTER(P), ctypes.c_int]
    lib.pdn_destroy.argtypes = [P]
    lib.pdn_strerror.argtypes = [ctypes.c_int]
    lib.pdn_strerror.restype = ctypes.c_char_p
    lib.pdn_last_error.argtypes = [P]
    lib.pdn_last_error.restype = ctypes.c_char_p
    lib.pdn_launch_count.argtypes = [P]
    lib.pdn_launch_count.restype = I64
    lib.pdn_stream.argtypes = [P]
    lib.pdn_stream.restype = ctypes.c_uint64
    lib.pdn_element_weights.argtypes = [P, I64, P, I64, P, P, ctypes.c_int]
    lib.pdn_assemble.argtypes = [P, I32, P, P, P, P, P, I64, I64, P, P, P, P, ctypes.c_int, ctypes.POINTER(I64)]
    lib.pdn_assemble_part.argtypes = [P, I32, P, P, P, P, P, I64, I64, I64, P, P, P, P, ctypes.c_int, ctypes.POINTER(I64)]
    lib.pdn_get_csr.argtypes = [P, P, P, P, ctypes.c_int]
    lib.pdn_get_shape.argtypes = [P, ctypes.POINTER(I64), ctypes.POINTER(I64)]
    lib.pdn_set_csr.argtypes = [P, I64, I64, P, P, P, ctypes.c_int]
    lib.pdn_solve.argtypes = [P, P, ctypes.POINTER(PdnOpts), P, ctypes.POINTER(PdnInfo), ctypes.c_int]
    lib.pdn_spmv.argtypes = [P, P, P, ctypes.c_int]
    lib.pdn_residual_noalloc.argtypes = [ctypes.c_int, ctypes.c_uint64, I64, P, P, P, P, P, P, P, ctypes.c_char_p, I64]
    lib.pdn_bench_spmv.argtypes = [P, I32, ctypes.c_int, ctypes.POINTER(D)]
    lib.pdn_power_density.argtypes = [P, I64, P, P, I64, P, D, P, ctypes.c_int]
    lib.pdn_nearest_vertex.argtypes = [P, I32, P, P, P, I64, P, P, P, P, ctypes.c_int]
    lib.pdn_format_f64.argtypes = [P, I64, P, I64]
    lib.pdn_format_f64.restype = I64
    lib.pdn_vtu_piece_bound.argtypes = [I64, I64]
    lib.pdn_vtu_piece_bound.restype = I64
    lib.pdn_vtu_piece.argtypes = [I64, P, P, I64, P, P, I64]
    lib.pdn_vtu_piece.restype = I64
    lib.pdn_comm_unique_id.argtypes = [P]
    lib.pdn_comm_init.argtypes = [P, ctypes.c_int, ctypes.c_int, P]
    lib.pdn_comm_destroy.argtypes = [P]
    lib.pdn_comm_info.argtypes = [P, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
    if path is None:
        _lib = lib
    return lib


def _raise(lib, ctx, status: int, what: str):
    detail = lib.pdn_last_error(ctx).decode() if ctx else ""
    text = lib.pdn_strerror(status).decode()
    msg = f"{what}: {text}" + (f" [{detail}]" if detail else "")
    if status == PDN_EINVAL:
        raise ValueError(msg)
    if status == PDN_EDEGENERATE:
        raise ZeroDivisionError("float division by zero" + (f" [{detail}]" if detail else ""))
    if status == PDN_ESINGULAR:
        raise SingularSystemError(msg)
    if status == PDN_ENOTCONVERGED:
        raise NotConvergedError(msg)
    if status == PDN_EUNSUPPORTED:
        raise NotImplementedError(msg)
    if status == PDN_ENONMANIFOLD:
        raise ValueError("Non-manifold mesh" + (f" [{detail}]" if detail else ""))
    if status == PDN_ENOMEM:
        raise MemoryError(msg)
    raise RuntimeError(msg)


def _ptr(a):
    """Host numpy array or device torch tensor -> void*."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(ctypes.c_void_p)
    return ctypes.c_void_p(a.data_ptr())        # torch tensor


def _is_device(a) -> bool:
    return not isinstance(a, np.ndarray) and hasattr(a, "is_cuda") and bool(a.is_cuda)


def _np(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


class Context:
    """One GPU, one stream, one resident system (``pdn_ctx``)."""

    def __init__(self, device: int = 0):
        self._lib = load_library()
        self._ctx = ctypes.c_void_p()
        st = self._lib.pdn_create(ctypes.byref(self._ctx), int(device))
        if st != PDN_OK:
            self._ctx = None
            raise RuntimeError(f"pdn_create(device={device}) failed: {self._lib.pdn_strerror(st).decode()} "
                               "(padne_b200 needs a CUDA device; there is no CPU fallback)")
        self.device = int(device)

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.pdn_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, st, what):
        if st != PDN_OK:
            _raise(self._lib, self._ctx, st, what)

    @property
    def launch_count(self) -> int:
        return int(self._lib.pdn_launch_count(self._ctx))

    @property
    def stream(self) -> int:
        return int(self._lib.pdn_stream(self._ctx))

    # -- multi-GPU ---------------------------------------------------------------------------------
    @staticmethod
    def comm_unique_id() -> bytes:
        """128-byte NCCL id (call on rank 0 and ship it to the other ranks)."""
        lib = load_library()
        buf = ctypes.create_string_buffer(128)
        st = lib.pdn_comm_unique_id(buf)
        if st != PDN_OK:
            raise RuntimeError(f"pdn_comm_unique_id: {lib.pdn_strerror(st).decode()}")
        return buf.raw

    def comm_init(self, rank: int, world: int, unique_id: bytes):
        """Collective: afterwards every :meth:`solve` on this context is a collective over ``world`` ranks."""
        if len(unique_id) != 128:
            raise ValueError("unique_id must be the 128 bytes of comm_unique_id()")
        buf = ctypes.create_string_buffer(bytes(unique_id), 128)
        self._check(self._lib.pdn_comm_init(self._ctx, int(rank), int(world), buf), "comm_init")

    def comm_destroy(self):
        self._check(self._lib.pdn_comm_destroy(self._ctx), "comm_destroy")

    @property
    def comm(self):
        """``(rank, world)`` of this context (``(0, 1)`` without a communicator)."""
        r, w = ctypes.c_int(), ctypes.c_int()
        self._lib.pdn_comm_info(self._ctx, ctypes.byref(r), ctypes.byref(w))
        return int(r.value), int(w.value)

    # -- kernel (1) ------------------------------------------------------------------------------
    def element_weights(self, xy, tri):
        """``[n_t, 3]`` corner weights ``abs(dot/cross)/2`` (``padne/mesh.py:124-139``)."""
        if _is_device(xy):
            import torch
            w = torch.empty((tri.shape[0], 3), dtype=torch.float64, device=xy.device)
            self._check(self._lib.pdn_element_weights(self._ctx, xy.shape[0], _ptr(xy), tri.shape[0], _ptr(tri),
                                                      _ptr(w), 1), "element_weights")
            return w
        xy = _np(xy, np.float64).reshape(-1, 2)
        tri = _np(tri, np.int32).reshape(-1, 3)
        w = np.empty((tri.shape[0], 3), np.float64)
        self._check(self._lib.pdn_element_weights(self._ctx, xy.shape[0], _ptr(xy), tri.shape[0], _ptr(tri),
                                                  _ptr(w), 0), "element_weights")
        return w

    # -- kernels (1)+(2) -----------------------------------------------------------------------------
    def assemble(self, v_off, xy, t_off, tri, conductance, N, s_row, s_col, s_val, s_op, n_nodes=None) -> int:
        """``assemble_system`` (``padne/solver.py:783-812``); the CSR stays resident.  Returns nnz.

        All arrays host (numpy) or all device (torch cuda tensors).  ``n_nodes`` (voltage rows) on a context
        with a communicator: partitioned assembly -- this rank builds (and, from host arrays, uploads) only its
        share (``pdn_assemble_part``)."""
        # whatever was resident is about to be replaced: solver.solve_system must not mistake it for an earlier
        # ResidentSystem (solver.assemble_flat sets the new token afterwards)
        self._resident_token = None
        dev = _is_device(xy)
        if not dev:
            v_off, t_off = _np(v_off, np.int64), _np(t_off, np.int64)
            xy, tri = _np(xy, np.float64), _np(tri, np.int32)
            conductance = _np(conductance, np.float64)
            s_row, s_col, s_op = _np(s_row, np.int32), _np(s_col, np.int32), _np(s_op, np.int32)
            s_val = _np(s_val, np.float64)
        nnz = ctypes.c_int64()
        if n_nodes is not None:
            st = self._lib.pdn_assemble_part(self._ctx, int(conductance.shape[0]), _ptr(v_off), _ptr(xy), _ptr(t_off),
                                             _ptr(tri), _ptr(conductance), int(N), int(n_nodes), int(s_row.shape[0]),
                                             _ptr(s_row), _ptr(s_col), _ptr(s_val), _ptr(s_op), int(dev), ctypes.byref(nnz))
            self._check(st, "assemble_part")
            return int(nnz.value)
        st = self._lib.pdn_assemble(self._ctx, int(conductance.shape[0]), _ptr(v_off), _ptr(xy), _ptr(t_off),
                                    _ptr(tri), _ptr(conductance), int(N), int(s_row.shape[0]), _ptr(s_row),
                                    _ptr(s_col), _ptr(s_val), _ptr(s_op), int(dev), ctypes.byref(nnz))
        self._check(st, "assemble")
        return int(nnz.value)

    def shape(self):
        N, nnz = ctypes.c_int64(), ctypes.c_int64()
        self._check(self._lib.pdn_get_shape(self._ctx, ctypes.byref(N), ctypes.byref(nnz)), "get_shape")
        return int(N.value), int(nnz.value)

    def get_csr(self):
        """``(indptr, indices, data)`` of the resident system as numpy arrays (``L.tocsr()``)."""
        N, nnz = self.shape()
        indptr = np.empty(N + 1, np.int32)
        indices = np.empty(nnz, np.int32)
        data = np.empty(nnz, np.float64)
        self._check(self._lib.pdn_get_csr(self._ctx, _ptr(indptr), _ptr(indices), _ptr(data), 0), "get_csr")
        return indptr, indices, data

    def set_csr(self, indptr, indices, data):
        self._resident_token = None
        dev = _is_device(data)
        if not dev:
            indptr, indices, data = _np(indptr, np.int32), _np(indices, np.int32), _np(data, np.float64)
        N = int(indptr.shape[0]) - 1
        self._check(self._lib.pdn_set_csr(self._ctx, N, int(data.shape[0]), _ptr(indptr), _ptr(indices),
                                          _ptr(data), int(dev)), "set_csr")

    # -- kernels (3)+(4) -----------------------------------------------------------------------------
    def solve(self, rhs, n_nodes: int = -1, rtol: float = 0.0, atol: float = 0.0, max_iter: int = 0,
              method: int = METHOD_AUTO, precond: int = PRECOND_AUTO, check_every: int = 0, out=None,
              n_coarse: int = 0):
        """``solve_system`` (``padne/solver.py:767-780``) on the resident system -> ``(v, SolveInfo)``."""
        dev = _is_device(rhs)
        N, _ = self.shape()
        if dev:
            import torch
            v = out if out is not None else torch.empty(N, dtype=torch.float64, device=rhs.device)
        else:
            rhs = _np(rhs, np.float64)
            v = out if out is not None else np.empty(N, np.float64)
        if rhs.shape[0] != N:
            raise ValueError(f"rhs has {rhs.shape[0]} entries, the system has {N}")
        opts = PdnOpts(rtol=rtol, atol=atol, max_iter=max_iter, method=method, precond=precond,
                       check_every=check_every, n_nodes=n_nodes, n_coarse=n_coarse)
        info = PdnInfo()
        st = self._lib.pdn_solve(self._ctx, _ptr(rhs), ctypes.byref(opts), _ptr(v), ctypes.byref(info), int(dev))
        self._check(st, "solve")
        return v, SolveInfo(METHOD_NAMES.get(info.method, "?"), info.iterations, info.restarts,
                            info.n_constraints, info.rel_residual, info.residual_norm,
                            info.ground_node_current, info.t_setup_ms, info.t_krylov_ms, info.t_total_ms,
                            PRECOND_NAMES.get(info.precond, "?"), info.n_coarse, info.t_coarse_ms,
                            info.n_coarse_local, info.local_level, info.stagnated, info.iteration_kind,
                            info.abs_target, info.floor_abs, info.rhs_norm, info.spmv_kind, info.spmv_halo,
                            info.spmv_window_coverage)

    def spmv(self, x):
        dev = _is_device(x)
        N, _ = self.shape()
        if dev:
            import torch
            y = torch.empty(N, dtype=torch.float64, device=x.device)
        else:
            x = _np(x, np.float64)
            y = np.empty(N, np.float64)
        self._check(self._lib.pdn_spmv(self._ctx, _ptr(x), _ptr(y), int(dev)), "spmv")
        return y

    def bench_spmv(self, reps: int = 50, mode: int = 0) -> float:
        """Mean ms per launch of the solver's SpMV kernel; mode 0 warm L2, 1 cold L2, 2 in situ
        (interleaved with the PCG vector update like a solver iteration)."""
        ms = ctypes.c_double()
        self._check(self._lib.pdn_bench_spmv(self._ctx, int(reps), int(mode), ctypes.byref(ms)), "bench_spmv")
        return float(ms.value)

    # -- row F1 ---------------------------------------------------------------------------------------
    def power_density(self, xy, values, tri, conductivity: float):
        """``compute_power_density`` (``padne/solver.py:728-745``) per face."""
        if _is_device(xy):
            import torch
            p = torch.empty(tri.shape[0], dtype=torch.float64, device=xy.device)
            self._check(self._lib.pdn_power_density(self._ctx, xy.shape[0], _ptr(xy), _ptr(values), tri.shape[0],
                                                    _ptr(tri), float(conductivity), _ptr(p), 1), "power_density")
            return p
        xy = _np(xy, np.float64).reshape(-1, 2)
        values = _np(values, np.float64)
        tri = _np(tri, np.int32).reshape(-1, 3)
        p = np.empty(tri.shape[0], np.float64)
        self._check(self._lib.pdn_power_density(self._ctx, xy.shape[0], _ptr(xy), _ptr(values), tri.shape[0],
                                                _ptr(tri), float(conductivity), _ptr(p), 0), "power_density")
        return p


    # -- row F3 ---------------------------------------------------------------------------------------
    def nearest_vertex(self, v_off, mesh_layer, xy, q_xy, q_layer, return_d2: bool = False):
        """Nearest vertex of every query point among the meshes of its layer: the ``KDTree.query`` loop of
        ``NodeIndexer.create`` (``padne/solver.py:421-433``).  Returns global vertex indices (int64, -1 when
        the layer is empty) and, optionally, the squared distances."""
        v_off = _np(v_off, np.int64)
        mesh_layer = _np(mesh_layer, np.int32)
        xy = _np(xy, np.float64).reshape(-1, 2)
        q_xy = _np(q_xy, np.float64).reshape(-1, 2)
        q_layer = _np(q_layer, np.int32)
        if v_off.shape[0] != mesh_layer.shape[0] + 1 or q_xy.shape[0] != q_layer.shape[0]:
            raise ValueError("nearest_vertex: inconsistent array lengths")
        if v_off[-1] != xy.shape[0]:
            raise ValueError("nearest_vertex: v_off[-1] must equal the number of vertices")
        out = np.empty(q_xy.shape[0], np.int64)
        d2 = np.empty(q_xy.shape[0], np.float64) if return_d2 else None
        self._check(self._lib.pdn_nearest_vertex(self._ctx, mesh_layer.shape[0], _ptr(v_off), _ptr(mesh_layer), _ptr(xy),
                                                 q_xy.shape[0], _ptr(q_xy), _ptr(q_layer), _ptr(out),
                                                 _ptr(d2) if return_d2 else None, 0), "nearest_vertex")
        return (out, d2) if return_d2 else out


# -- row F4: host-only text formatting (no context, no GPU) -------------------------------------------
def format_f64(x) -> bytes:
    """``" ".join(str(v) for v in x)`` for a float64 array (``padne/paraview.py:91``), CPython's repr."""
    lib = load_library()
    x = _np(x, np.float64).reshape(-1)
    buf = np.empty(max(1, 25 * x.shape[0]), np.uint8)
    n = lib.pdn_format_f64(_ptr(x), x.shape[0], _ptr(buf), buf.shape[0])
    if n < 0:
        _raise(lib, None, int(-n), "format_f64")
    return buf[:n].tobytes()


def vtu_piece(xy, tri, voltage) -> np.ndarray:
    """The ``<Piece>`` element of one mesh (``padne/paraview.py:209-231``) as a uint8 array of XML text."""
    lib = load_library()
    xy = _np(xy, np.float64).reshape(-1, 2)
    tri = _np(tri, np.int32).reshape(-1, 3)
    voltage = _np(voltage, np.float64).reshape(-1)
    if voltage.shape[0] != xy.shape[0]:
        raise ValueError("vtu_piece: one voltage per vertex")
    buf = np.empty(lib.pdn_vtu_piece_bound(xy.shape[0], tri.shape[0]), np.uint8)
    n = lib.pdn_vtu_piece(xy.shape[0], _ptr(xy), _ptr(voltage), tri.shape[0], _ptr(tri), _ptr(buf), buf.shape[0])
    if n < 0:
        _raise(lib, None, int(-n), "vtu_piece")
    return buf[:n]


_default_ctx: dict[int, Context] = {}


NOALLOC_MAX_ROWS = 1 << 24


def residual_noalloc(indptr, indices, data, x, b, r, rr, device: int = 0, stream: int = 0, n_rows: int | None = None) -> None:
    """``pdn_residual_noalloc``: r = b - A x and rr[0] = ||r||^2 into CALLER-OWNED device tensors, with no device
    allocation anywhere on the path (no reference counterpart; see include/padne_b200.h).  Asynchronous on
    ``stream`` (a ``cudaStream_t`` as an integer; 0 = the default stream).  At most ``NOALLOC_MAX_ROWS`` rows."""
    lib = load_library()
    n = int(indptr.shape[0]) - 1 if n_rows is None else int(n_rows)
    err = ctypes.create_string_buffer(512)
    st = lib.pdn_residual_noalloc(int(device), int(stream), n, _ptr(indptr), _ptr(indices), _ptr(data), _ptr(x),
                                  _ptr(b), _ptr(r), _ptr(rr), err, len(err))
    if st == PDN_OK:
        return
    msg = err.value.decode() or f"pdn_residual_noalloc: {lib.pdn_strerror(st).decode()}"
    if st == PDN_EINVAL:
        raise ValueError(msg)
    if st == PDN_EUNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)


def default_context(device: int = 0) -> Context:
    """Process-wide context per device (created on first use)."""
    ctx = _default_ctx.get(device)
    if ctx is None or ctx._ctx is None:
        ctx = Context(device)
        _default_ctx[device] = ctx
    return ctx


__all__ = ["Context", "default_context", "load_library", "SolveInfo", "SingularSystemError",
           "NotConvergedError", "SYMBOLS", "LIB_PATH", "warnings", "residual_noalloc", "NOALLOC_MAX_ROWS"]
