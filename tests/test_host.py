"""Host logic and the C-ABI surface -- CPU only (no compute calls)."""
import ctypes
import os
import re

import numpy as np
import pytest
import scipy.sparse

from conftest import ROOT, golden_system_files, load_system
from padne_b200 import capi, synth
from padne_b200.flat import FlatSystem, KIND_REGULATOR, KIND_VOLTAGE, OP_ADD, OP_SET
from padne_b200.stamps import build_rhs, build_stamps


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "padne_b200.h")).read()
    declared = set(re.findall(r"\b(pdn_[a-z0-9_]+)\s*\(", header))
    declared -= {"pdn_ctx"}          # "pdn_ctx (the reference ..." in a comment
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    lib = capi.load_library()
    for name in declared:
        assert hasattr(lib, name), name


def test_strerror_and_struct_layout():
    lib = capi.load_library()
    assert lib.pdn_strerror(0) == b"ok"
    assert b"singular" in lib.pdn_strerror(capi.PDN_ESINGULAR)
    assert lib.pdn_strerror(capi.PDN_ENONMANIFOLD) == b"Non-manifold mesh"
    assert ctypes.sizeof(capi.PdnOpts) == 48
    assert ctypes.sizeof(capi.PdnInfo) == 144
    # the ctypes mirrors against the header itself: gcc prints the C sizes and the offset of the last field
    import subprocess
    import tempfile
    src = ('#include <stdio.h>\n#include <stddef.h>\n#include "padne_b200.h"\n'
           'int main(void){printf("%zu %zu %zu", sizeof(pdn_opts), sizeof(pdn_info), '
           'offsetof(pdn_info, spmv_window_coverage));return 0;}\n')
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        so, si, off = map(int, subprocess.check_output([exe]).split())
    assert (so, si) == (ctypes.sizeof(capi.PdnOpts), ctypes.sizeof(capi.PdnInfo))
    assert off == capi.PdnInfo.spmv_window_coverage.offset


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        capi.Context(0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "padne_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "padne_oracle" not in src, f


def test_stamp_order_and_ops():
    z = np.zeros(1, np.int64)
    fs = FlatSystem(v_off=z, t_off=z, xy=np.zeros((0, 2)), tri=np.zeros((0, 3), np.int32), conductance=np.zeros(0),
                    n_internal=4, n_extra=2, i_gnd=1,
                    e_kind=[0, 2, 3, 1], e_node=[[0, 1, 0, 0], [1, 2, 0, 0], [2, 3, 0, 1], [3, 0, 0, 0]],
                    e_val=[2.0, 5.0, 3.3, 0.25], e_gain=[0, 0, 0.5, 0], e_xidx=[-1, 4, 5, -1])
    s_row, s_col, s_val, s_op, r = build_stamps(fs)
    # resistor: 4 adds; source: 4 sets; regulator: 4 sets + 2 adds; ground: 2 sets
    assert list(s_op) == [OP_ADD] * 4 + [OP_SET] * 4 + [OP_SET] * 4 + [OP_ADD] * 2 + [OP_SET] * 2
    assert list(zip(s_row[:4], s_col[:4])) == [(0, 0), (0, 1), (1, 1), (1, 0)]
    assert list(s_val[:4]) == [-0.5, 0.5, -0.5, 0.5]
    assert list(zip(s_row[4:8], s_col[4:8], s_val[4:8])) == [(4, 1, 1.0), (4, 2, -1.0), (1, 4, 1.0), (2, 4, -1.0)]
    assert list(zip(s_row[12:14], s_col[12:14], s_val[12:14])) == [(0, 5, 0.5), (1, 5, -0.5)]
    assert list(zip(s_row[14:], s_col[14:])) == [(6, 1), (1, 6)]
    assert r[4] == 5.0 and r[5] == 3.3 and r[3] == 0.25 and r[0] == -0.25 and r[6] == 0.0


@pytest.mark.parametrize("path", golden_system_files(), ids=os.path.basename)
def test_rhs_matches_reference(path):
    fs, d = load_system(path)
    assert np.array_equal(build_rhs(fs), d["r"])


def test_synth_to_flat_matches_golden_inputs():
    # the generator is deterministic: regenerating a golden board gives the stored geometry
    fs, d = load_system([p for p in golden_system_files() if "2layer_24" in p][0])
    b = synth.board_2layer(24, n_via=2, size_mm=10.0, seed=2).to_flat()
    # the reference's face iteration starts at the LAST half-edge (mesh.py:323): (v3, v1, v2)
    assert np.array_equal(b.xy, fs.xy) and np.array_equal(np.roll(b.tri, 1, axis=1), fs.tri)
    assert b.N == fs.N and b.n_elements == fs.n_elements
    assert np.array_equal(b.e_kind, fs.e_kind) and np.array_equal(b.e_val, fs.e_val)


def test_synth_sizes_of_the_baseline_configs():
    # 2 x 708^2 = 1 002 528 grid vertices before via holes
    b = synth.board_2layer(64, n_via=2)
    fs = b.to_flat()
    assert fs.n_mesh == 2 and fs.n_vertices == 2 * (64 * 64 - 4 * 9)
    assert fs.has_regulator is False
    r = synth.board_4layer_regulator(40, n_via=2).to_flat()
    assert r.has_regulator and r.n_extra == 3 and int((r.e_kind == KIND_REGULATOR).sum()) == 2
    assert int((r.e_kind == KIND_VOLTAGE).sum()) == 1


def test_flat_system_validation():
    with pytest.raises(ValueError):
        FlatSystem(v_off=[0, 3], t_off=[0, 1], xy=np.zeros((2, 2)), tri=np.zeros((1, 3), np.int32), conductance=[1.0])


def test_stamp_network_into_system_host_mirror():
    # same writes as the reference on a lil_matrix (tests/test_solver.py:21-63 style)
    from padne_b200 import solver as S

    class NodeID:
        pass

    class Resistor:
        def __init__(self, a, b, resistance):
            self.a, self.b, self.resistance = a, b, resistance

    class VoltageSource:
        def __init__(self, p, n, voltage):
            self.p, self.n, self.voltage = p, n, voltage

    class Net:
        pass

    class Idx:
        pass

    a, b = NodeID(), NodeID()
    vs = VoltageSource(b, a, 3.3)
    net = Net()
    net.elements = [vs, Resistor(a, b, 2.2)]
    idx = Idx()
    idx.node_to_global_index = {a: 0, b: 1}
    idx.extra_source_to_global_index = {vs: 2}
    L = scipy.sparse.lil_matrix((3, 3))
    r = np.zeros(3)
    S.stamp_network_into_system(net, idx, L, r)
    g = 1 / 2.2
    expect = np.array([[-g, g, -1.0], [g, -g, 1.0], [-1.0, 1.0, 0.0]])
    assert np.array_equal(L.toarray(), expect)
    assert np.array_equal(r, [0.0, 0.0, 3.3])


def test_node_indexer_create_with_batched_lookup():
    # Row F3 drop-in: the reference's own NodeIndexer.create, its KD-trees swapped for one batched lookup
    # (the oracle stands in for the GPU here), must number every node exactly like the stock create().
    # Needs the reference importable (build container only); nothing is read from it on the GPU box.
    from oracle import refload
    if not refload.reference_available():
        pytest.skip("reference not present")
    from oracle import oracle
    from padne_b200 import solver as S
    problem, mesh, ref = refload.load_reference()
    import shapely.geometry as geometry
    for board in (synth.board_4layer_regulator(20, n_via=2, size_mm=8.0, seed=4),
                  synth.board_via_tht_standin(20, seed=1, n_via=2)):
        prob, meshes, m2l = synth.to_reference_objects(board, problem, mesh, geometry)
        vindex = ref.VertexIndexer.create(meshes)
        nets = list(prob.networks)
        stock = ref.NodeIndexer.create(prob, meshes, m2l, vindex, nets)
        calls = []

        def lookup(*a):
            calls.append(len(a[3]))
            return oracle.nearest_vertex(*a)[0]

        ours = S.node_indexer_create(ref, prob, meshes, m2l, vindex, nets, lookup=lookup)
        assert calls == [sum(len(n.connections) for n in nets)]           # ONE batched call
        assert ours.internal_node_count == stock.internal_node_count
        assert ours.node_to_global_index == stock.node_to_global_index    # same NodeID objects, same indices
        assert ours.extra_source_to_global_index == stock.extra_source_to_global_index
        assert "_construct_kdtrees" in ref.NodeIndexer.__dict__            # the reference class is restored
        again = ref.NodeIndexer.create(prob, meshes, m2l, vindex, nets)
        assert again.node_to_global_index == stock.node_to_global_index


def test_solve_system_warns_only_when_the_requested_rtol_is_missed():
    # ADVICE (round 1): a caller asking for rtol must be told when the iteration stopped above it.  The C ABI accepts
    # stagnation only at the fp64 floor (pdn_info.stagnated); the Python mirror turns "stagnated AND above rtol" into
    # a SolverWarning and stays silent otherwise, like the reference's direct solve (padne/solver.py:767-780).
    import dataclasses
    import warnings
    from oracle_context import OracleContext
    from padne_b200 import solver as S

    class Stagnating(OracleContext):
        def __init__(self, rel, stagnated):
            super().__init__()
            self.rel, self.stag = rel, stagnated

        def solve(self, rhs, n_nodes=-1, **opts):
            v, info = super().solve(rhs, n_nodes, **opts)
            return v, dataclasses.replace(info, rel_residual=self.rel, stagnated=self.stag, floor_abs=1e-9, rhs_norm=2e3)

    fs, d = load_system(golden_system_files()[0])
    L = scipy.sparse.csr_matrix((d["data"], d["indices"], d["indptr"]), shape=(fs.N, fs.N))
    for rel, stag, expect in ((3e-10, 1, True), (5e-11, 1, False), (3e-10, 0, False)):
        with warnings.catch_warnings(record=True) as rec:
            warnings.simplefilter("always")
            v, info = S.solve_system(L, d["r"], ctx=Stagnating(rel, stag), n_nodes=fs.n_nodes)
        hit = [w for w in rec if issubclass(w.category, S.SolverWarning) and "stagnated" in str(w.message)]
        assert bool(hit) == expect, (rel, stag, [str(w.message) for w in rec])
        assert np.allclose(v, d["v"], rtol=1e-9, atol=1e-12) and info.residual_norm < 1e-9
