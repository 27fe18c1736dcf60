"""A stand-in for ``padne_b200.capi.Context`` whose arithmetic is the CPU oracle.  TEST INFRASTRUCTURE ONLY.

The host side of the drop-in (``padne_b200/solver.py``: flattening, stamp order, RHS, the reference's classes in the
returned ``Solution``, warnings) has to be checked where there is no GPU.  This object has the method signatures of
``capi.Context`` that ``padne_b200.solver`` calls and answers them with ``oracle/``, so the CPU suite can run
``solve_meshed`` end to end against the unmodified reference.  Nothing under ``padne_b200/`` knows it exists.
"""
import types

import numpy as np
import scipy.sparse
import scipy.sparse.linalg

from oracle import oracle
from padne_b200 import capi


class OracleContext:
    def __init__(self):
        self.csr = None
        self.calls = []

    def assemble(self, v_off, xy, t_off, tri, conductance, N, s_row, s_col, s_val, s_op) -> int:
        fs = types.SimpleNamespace(n_mesh=len(conductance), v_off=np.ascontiguousarray(v_off, np.int64),
                                   xy=np.ascontiguousarray(xy, np.float64), t_off=np.ascontiguousarray(t_off, np.int64),
                                   tri=np.ascontiguousarray(tri, np.int32),
                                   conductance=np.ascontiguousarray(conductance, np.float64), N=int(N))
        self.csr = oracle.assemble(fs, (np.asarray(s_row, np.int32), np.asarray(s_col, np.int32),
                                        np.asarray(s_val, np.float64), np.asarray(s_op, np.int32)))
        self.calls.append("assemble")
        return int(self.csr[2].shape[0])

    def shape(self):
        return len(self.csr[0]) - 1, int(self.csr[2].shape[0])

    def get_csr(self):
        return self.csr

    def set_csr(self, indptr, indices, data):
        self.csr = (np.asarray(indptr, np.int32), np.asarray(indices, np.int32), np.asarray(data, np.float64))
        self.calls.append("set_csr")

    def solve(self, rhs, n_nodes=-1, **opts):
        self.calls.append("solve")
        indptr, indices, data = self.csr
        n = len(indptr) - 1
        A = scipy.sparse.csr_matrix((data, indices, indptr), shape=(n, n))
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("error")
            try:
                v = scipy.sparse.linalg.spsolve(A.tocsc(), rhs)
            except scipy.sparse.linalg.MatrixRankWarning as e:
                raise capi.SingularSystemError(str(e))
        res = float(np.linalg.norm(A @ v - rhs))
        return v, capi.SolveInfo("direct", 0, 0, 0, res / max(float(np.linalg.norm(rhs)), 1e-300), res, float(v[-1]),
                                 0.0, 0.0, 0.0)

    def power_density(self, xy, values, tri, conductivity):
        self.calls.append("power_density")
        return oracle.power_density(xy, values, tri, conductivity)

    def nearest_vertex(self, v_off, mesh_layer, xy, q_xy, q_layer, return_d2=False):
        self.calls.append("nearest_vertex")
        idx, d2 = oracle.nearest_vertex(v_off, mesh_layer, xy, q_xy, q_layer)
        return (idx, d2) if return_d2 else idx
