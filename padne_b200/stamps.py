"""Coupling rows: lumped elements -> ordered matrix stamp operations and the RHS.

Host-side mirror of ``stamp_network_into_system`` (``padne/solver.py:469-541``) and
``setup_ground_node`` (``solver.py:544-560``).  The reference writes each stamp straight into a
``scipy.sparse.lil_matrix``; here the same writes are emitted, in the same order, as an
operation list ``(row, col, value, op)`` that assembly kernel (2) replays on the device with LIL
semantics:

* ``op == OP_ADD`` is ``L[row, col] += value`` (resistor and regulator-gain stamps);
* ``op == OP_SET`` is ``L[row, col] = value`` (voltage-source / ground stamps overwrite);
* an entry whose value ends up exactly ``0.0`` is absent from the pattern (LIL deletes it).

All arithmetic that the reference does in Python floats (``1 / resistance``, ``-gain``) is done
here in IEEE fp64 element-wise numpy, which rounds identically.
"""
from __future__ import annotations

import numpy as np

from .flat import (FlatSystem, KIND_CURRENT, KIND_REGULATOR, KIND_RESISTOR, KIND_VOLTAGE,
                   OP_ADD, OP_SET)

_SLOTS = 8  # op slots reserved per element when merging the per-kind tables back into order


def build_stamps(fs: FlatSystem, with_ground: bool = True):
    """Return ``(s_row, s_col, s_val, s_op, rhs)`` for ``fs``.

    ``s_*`` are in the exact order the reference would perform the LIL writes
    (element order, then the per-element order of ``solver.py:475-538``, then the ground stamp).
    ``rhs`` is the vector ``r`` of ``solver.py:763`` after all stamps.
    """
    N = fs.N
    kind, node, val, gain, xidx = fs.e_kind, fs.e_node, fs.e_val, fs.e_gain, fs.e_xidx
    eidx = np.arange(fs.n_elements, dtype=np.int64)
    rows, cols, vals, ops, keys = [], [], [], [], []

    def emit(sel, r, c, v, op, slot):
        rows.append(r.astype(np.int32))
        cols.append(c.astype(np.int32))
        vals.append(np.asarray(v, dtype=np.float64))
        ops.append(np.full(r.shape[0], op, np.int32))
        keys.append(eidx[sel] * _SLOTS + slot)

    # Resistor (solver.py:475-484)
    sel = np.nonzero(kind == KIND_RESISTOR)[0]
    if sel.size:
        a, b = node[sel, 0], node[sel, 1]
        g = 1.0 / val[sel]
        emit(sel, a, a, -g, OP_ADD, 0)
        emit(sel, a, b, g, OP_ADD, 1)
        emit(sel, b, b, -g, OP_ADD, 2)
        emit(sel, b, a, g, OP_ADD, 3)
    # VoltageSource (solver.py:493-511) and the source half of VoltageRegulator (solver.py:526-529)
    for k in (KIND_VOLTAGE, KIND_REGULATOR):
        sel = np.nonzero(kind == k)[0]
        if not sel.size:
            continue
        p, n, iv = node[sel, 0], node[sel, 1], xidx[sel]
        one = np.ones(sel.size)
        emit(sel, iv, p, one, OP_SET, 0)
        emit(sel, iv, n, -one, OP_SET, 1)
        emit(sel, p, iv, one, OP_SET, 2)
        emit(sel, n, iv, -one, OP_SET, 3)
        if k == KIND_REGULATOR:  # current mirror, non-symmetric (solver.py:537-538)
            sf, st = node[sel, 2], node[sel, 3]
            emit(sel, sf, iv, gain[sel], OP_ADD, 4)
            emit(sel, st, iv, -gain[sel], OP_ADD, 5)

    if rows:
        key = np.concatenate(keys)
        order = np.argsort(key, kind="stable")
        s_row = np.concatenate(rows)[order]
        s_col = np.concatenate(cols)[order]
        s_val = np.concatenate(vals)[order]
        s_op = np.concatenate(ops)[order]
    else:
        s_row = np.zeros(0, np.int32)
        s_col = np.zeros(0, np.int32)
        s_val = np.zeros(0, np.float64)
        s_op = np.zeros(0, np.int32)

    if with_ground:  # solver.py:558-560
        s_row = np.concatenate([s_row, np.array([N - 1, fs.i_gnd], np.int32)])
        s_col = np.concatenate([s_col, np.array([fs.i_gnd, N - 1], np.int32)])
        s_val = np.concatenate([s_val, np.array([1.0, 1.0])])
        s_op = np.concatenate([s_op, np.array([OP_SET, OP_SET], np.int32)])

    rhs = build_rhs(fs)
    return (np.ascontiguousarray(s_row), np.ascontiguousarray(s_col),
            np.ascontiguousarray(s_val), np.ascontiguousarray(s_op), rhs)


def build_rhs(fs: FlatSystem) -> np.ndarray:
    """RHS ``r`` (``solver.py:489-492, 505, 530, 560``)."""
    r = np.zeros(fs.N, dtype=np.float64)
    kind, node, val, xidx = fs.e_kind, fs.e_node, fs.e_val, fs.e_xidx
    sel = np.nonzero(kind == KIND_CURRENT)[0]
    if sel.size:
        idx = np.empty(2 * sel.size, np.int64)
        inc = np.empty(2 * sel.size, np.float64)
        idx[0::2], idx[1::2] = node[sel, 0], node[sel, 1]
        inc[0::2], inc[1::2] = val[sel], -val[sel]
        np.add.at(r, idx, inc)  # unbuffered, in order: same rounding as the sequential +=
    sel = np.nonzero(kind == KIND_VOLTAGE)[0]
    if sel.size:
        r[xidx[sel]] = val[sel]           # assignment: the later element wins
    sel = np.nonzero(kind == KIND_REGULATOR)[0]
    if sel.size:
        np.add.at(r, xidx[sel], val[sel])  # r[i_v] += voltage
    # r[-1] = 0 for the ground row is already the case
    return r
