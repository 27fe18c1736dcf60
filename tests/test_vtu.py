"""Row F4: ``padne_b200.paraview`` / ``pdn_vtu_piece`` -- CPU only, host code.

Pinned by ``tests/golden/vtu_*.json.gz``: the element trees the reference's own ``export_solution``
(``padne/paraview.py:234-298``) builds for reference ``Solution`` objects (generator:
``tests/golden/make_golden_vtu.py``).  The filename cases restate ``tests/test_paraview.py:16-66`` of the
reference, the structural ones ``:138-146, 348-405, 517-600``.
"""
import glob
import gzip
import json
import os
import types
import xml.etree.ElementTree as ET

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import oracle
from padne_b200 import capi, paraview, synth
from padne_b200.flat import FlatSystem


def _tree(e):
    """Parsed element -> the golden's list form; whitespace-only text (pretty-printing) counts as none."""
    text = e.text if e.text is not None and e.text.strip() != "" else None
    return [e.tag, [[k, v] for k, v in e.attrib.items()], text, [_tree(c) for c in e]]


def _norm(tree):
    """The reference sets ``text = ""`` on an empty DataArray; a parser reads that back as no text."""
    tag, attrs, text, kids = tree
    return [tag, attrs, text or None, [_norm(k) for k in kids]]


# ---------------------------------------------------------------------------------------------- str(float)
EDGE = [0.0, -0.0, 1.0, -1.0, 1.5, 0.1, 0.1 + 0.2, 1e-4, 1e-5, 0.00012345, 9.999e-5, 123.0, 1e15, 1e16,
        9999999999999998.0, 123456789012345678.0, 1e22, 1e23, 1e100, 1e-100, 5e-324, 2.2250738585072014e-308,
        1.7976931348623157e308, -1.7976931348623157e308, 9007199254740993.0, float("inf"), float("-inf"),
        float("nan"), 1 / 3, 2 / 3, 100.0, 1e-7, 123456.789, 0.5, 2.0 ** 53, 2.0 ** -30]


def test_format_matches_cpython_repr_on_edge_values():
    for x in EDGE:
        assert capi.format_f64([x]).decode() == repr(x)
    assert capi.format_f64(EDGE).decode() == " ".join(repr(x) for x in EDGE)
    assert capi.format_f64([]) == b""


@pytest.mark.parametrize("n", [1000, 200_000])          # serial path / chunked parallel path
def test_format_matches_cpython_repr_on_random_bit_patterns(n):
    rng = np.random.default_rng(n)
    a = rng.integers(0, 2 ** 64, size=n, dtype=np.uint64).view(np.float64)
    assert capi.format_f64(a).decode() == " ".join(map(repr, a.tolist()))
    b = np.round(rng.random(n) * 10.0 ** rng.integers(-8, 20, n).astype(np.float64), 3)   # board-like values
    assert capi.format_f64(b).decode() == " ".join(map(repr, b.tolist()))
    assert capi.format_f64(b).decode() == " ".join(str(np.float64(x)) for x in b[:500]) + \
        (" " + " ".join(map(repr, b[500:].tolist())) if n > 500 else "")


def test_format_around_every_power_of_ten_and_two():
    """The notation switches at 1e-4 and 1e16 and the digit count changes at every power of ten: the values on
    both sides of each, and of each power of two (where the spacing of doubles changes), subnormals included."""
    xs = []
    for e in range(-324, 309):
        x = float(f"1e{e}")
        xs += [np.nextafter(x, 0.0), x, np.nextafter(x, np.inf)]
    for e in range(-1074, 1024):
        x = float(np.ldexp(1.0, e))
        xs += [np.nextafter(x, 0.0), x, np.nextafter(x, np.inf)]
    xs = np.array(xs + [-x for x in xs], np.float64)
    assert capi.format_f64(xs).decode() == " ".join(map(repr, xs.tolist()))


def test_thread_count_does_not_change_the_text(monkeypatch):
    a = np.random.default_rng(5).normal(size=150_001) * 1e3
    want = " ".join(map(repr, a.tolist())).encode()
    for t in ("1", "3", "16"):
        monkeypatch.setenv("PADNE_B200_IO_THREADS", t)
        assert capi.format_f64(a) == want


# ------------------------------------------------------------------------------------------------ filenames
@pytest.mark.parametrize("name, stem", [("F.Cu", "F.Cu"), ("Power Plane", "Power_Plane"), ("Layer/Name?!@", "Layer_Name"),
                                        ("", "layer"), ("   ", "layer"), ("Test___Name", "Test_Name"),
                                        ("In1-Cu.2", "In1-Cu.2"), ("???", "layer"), ("_x_", "x")])
def test_sanitize_filename(name, stem):
    used = set()
    assert paraview._sanitize_filename(name, used) == stem
    assert used == {stem}


def test_sanitize_filename_duplicates():
    used = set()
    assert [paraview._sanitize_filename("F.Cu", used) for _ in range(3)] == ["F.Cu", "F.Cu_2", "F.Cu_3"]
    assert used == {"F.Cu", "F.Cu_2", "F.Cu_3"}
    assert paraview._sanitize_filename("", used, fallback_prefix="mesh") == "mesh"


# ------------------------------------------------------------------------------------------- one piece, bytes
UNIT_SQUARE = b"""<?xml version='1.0' encoding='UTF-8'?>
<VTKFile type="UnstructuredGrid" version="0.1" byte_order="LittleEndian">
  <UnstructuredGrid>
    <Piece NumberOfPoints="5" NumberOfCells="4">
      <PointData Scalars="voltage">
        <DataArray type="Float64" format="ascii" Name="voltage">0.0 1.0 3.3 1e-05 0.5</DataArray>
      </PointData>
      <Points>
        <DataArray type="Float64" format="ascii" NumberOfComponents="3">0.0 -0.0 0.0 1.0 -0.0 0.0 1.0 -1.0 0.0 0.0 -1.0 0.0 0.5 -0.5 0.0</DataArray>
      </Points>
      <Cells>
        <DataArray type="Int32" format="ascii" Name="connectivity">0 1 4 1 2 4 2 3 4 3 0 4</DataArray>
        <DataArray type="Int32" format="ascii" Name="offsets">3 6 9 12</DataArray>
        <DataArray type="UInt8" format="ascii" Name="types">5 5 5 5</DataArray>
      </Cells>
    </Piece>
  </UnstructuredGrid>
</VTKFile>
"""


def test_unit_square_document_bytes(tmp_path):
    """The whole file for the unit square + centre of ``tests/test_solver.py:776-852``."""
    xy = [(0, 0), (1, 0), (1, 1), (0, 1), (0.5, 0.5)]
    tri = [(0, 1, 4), (1, 2, 4), (2, 3, 4), (3, 0, 4)]
    n = paraview.write_vtu(tmp_path / "a.vtu", [(xy, tri, [0.0, 1.0, 3.3, 1e-5, 0.5])])
    data = (tmp_path / "a.vtu").read_bytes()
    assert data == UNIT_SQUARE and n == len(data)
    root = ET.fromstring(data)
    assert _tree(root) == oracle.vtu_file_tree([(xy, tri, [0.0, 1.0, 3.3, 1e-5, 0.5])])


def test_empty_and_faceless_pieces():
    for xy, tri, v in [(np.zeros((0, 2)), np.zeros((0, 3), np.int32), np.zeros(0)),
                       ([(0.0, 0.0), (2.0, 1.0)], np.zeros((0, 3), np.int32), [1.0, 2.0])]:
        text = bytes(paraview.piece_text(xy, tri, v))
        assert _tree(ET.fromstring(text)) == _norm(oracle.vtu_piece_tree(xy, tri, v))
        assert b'Name="types"></DataArray>' in text


def test_bad_arguments():
    with pytest.raises(ValueError):
        paraview.piece_text([(0, 0), (1, 0), (0, 1)], [(0, 1, 3)], [0, 0, 0])       # index out of range
    with pytest.raises(ValueError):
        paraview.piece_text([(0, 0), (1, 0), (0, 1)], [(0, 1, -1)], [0, 0, 0])
    with pytest.raises(ValueError):
        paraview.piece_text([(0, 0), (1, 0), (0, 1)], [(0, 1, 2)], [0, 0])          # one voltage per vertex
    with pytest.raises(ValueError):
        paraview.write_vtu("/nonexistent/never_written.vtu", [])


@pytest.mark.parametrize("seed", [0, 1])
def test_piece_against_oracle_on_jittered_layers(seed):
    n = 190 if seed else 40                     # 190 x 190 vertices takes the parallel path
    xy, tri, _ = synth.grid_layer(n, n, 0.37, seed=seed, jitter=0.3)
    v = np.random.default_rng(seed).normal(size=len(xy)) * 10.0 ** np.random.default_rng(seed + 9).integers(-7, 3, len(xy))
    got = _tree(ET.fromstring(bytes(paraview.piece_text(xy, tri, v))))
    assert got == oracle.vtu_piece_tree(xy, tri, v)


# ------------------------------------------------------------------------------- golden: the reference's trees
def _golden_cases():
    return sorted(glob.glob(os.path.join(GOLDEN, "vtu_*.json.gz")))


def test_golden_present():
    assert len(_golden_cases()) >= 3


@pytest.mark.parametrize("path", _golden_cases(), ids=lambda p: os.path.basename(p)[4:-8])
def test_export_arrays_against_reference_trees(path, tmp_path):
    with gzip.open(path, "rt") as f:
        g = json.load(f)
    case = os.path.basename(path)[4:-8]
    d = np.load(os.path.join(GOLDEN, f"system_{case}.npz"))
    fs = FlatSystem.from_npz_dict(d, name=case)
    assert g["write_kwargs"] == {"xml_declaration": True, "encoding": "utf-8", "pretty_print": True}
    written = paraview.export_arrays(g["layer_names"], g["mesh_layer"], fs.v_off, fs.t_off, fs.xy, fs.tri, d["v"], tmp_path)
    assert sorted(p.name for p in written) == sorted(g["files"])
    assert sorted(os.listdir(tmp_path)) == sorted(g["files"])
    used, layer_of_file = set(), {}
    for li, name in enumerate(g["layer_names"]):
        if li in g["mesh_layer"]:
            layer_of_file[paraview._sanitize_filename(name, used) + ".vtu"] = li
    for p in written:
        data = p.read_bytes()
        assert data.startswith(b"<?xml version='1.0' encoding='UTF-8'?>\n<VTKFile ") and data.endswith(b"</VTKFile>\n")
        assert _tree(ET.fromstring(data)) == _norm(g["files"][p.name])
        # and the oracle restatement says the same
        members = [m for m, l in enumerate(g["mesh_layer"]) if l == layer_of_file[p.name]]
        pieces = [(fs.xy[fs.v_off[m]:fs.v_off[m + 1]], fs.tri[fs.t_off[m]:fs.t_off[m + 1]],
                   d["v"][fs.v_off[m]:fs.v_off[m + 1]]) for m in members]
        assert oracle.vtu_file_tree(pieces) == g["files"][p.name]


# ------------------------------------------------------------------------------ export_solution, drop-in form
def test_export_solution_on_solution_shaped_objects(tmp_path):
    """Layer naming, duplicate names, skipped empty layers, several meshes per layer (``:262-290``)."""
    def mesh_and_form(n, seed):
        xy, tri, _ = synth.grid_layer(n, n + 1, 0.5, seed=seed, jitter=0.2)
        msh = types.SimpleNamespace(xy=xy, tri=tri)
        return msh, types.SimpleNamespace(mesh=msh, values=np.random.default_rng(seed).random(len(xy)))

    (m0, f0), (m1, f1), (m2, f2) = mesh_and_form(4, 0), mesh_and_form(5, 1), mesh_and_form(3, 2)
    layers = [types.SimpleNamespace(name=n) for n in ("F.Cu", "Empty", "F.Cu", "Power Plane")]
    ls = [types.SimpleNamespace(meshes=[m0, m1], potentials=[f0, f1]),
          types.SimpleNamespace(meshes=[], potentials=[]),
          types.SimpleNamespace(meshes=[m2], potentials=[f2]),
          types.SimpleNamespace(meshes=[m1], potentials=[f1])]
    solution = types.SimpleNamespace(problem=types.SimpleNamespace(layers=layers), layer_solutions=ls)
    assert paraview.export_solution(solution, tmp_path / "out") is None
    assert sorted(os.listdir(tmp_path / "out")) == ["F.Cu.vtu", "F.Cu_2.vtu", "Power_Plane.vtu"]
    root = ET.fromstring((tmp_path / "out" / "F.Cu.vtu").read_bytes())
    assert _tree(root) == oracle.vtu_file_tree([(m0.xy, m0.tri, f0.values), (m1.xy, m1.tri, f1.values)])
    assert root.tag == "VTKFile" and root.get("type") == "UnstructuredGrid" and root.get("version") == "0.1"
    pieces = root.find("UnstructuredGrid").findall("Piece")
    assert [p.get("NumberOfPoints") for p in pieces] == [str(len(m0.xy)), str(len(m1.xy))]
    conn = np.array(pieces[0].find("Cells")[0].text.split(), np.int64)
    assert conn.size == 3 * int(pieces[0].get("NumberOfCells")) and conn.max() < len(m0.xy)


def test_export_solution_with_reference_objects(tmp_path):
    """Reference ``Mesh`` / ``ZeroForm`` / ``Solution`` objects in, the reference exporter's tree out."""
    from oracle import refload
    if not refload.reference_available():
        pytest.skip("reference not present")
    problem, mesh, solver = refload.load_reference()
    import shapely.geometry as geometry
    board = synth.board_2layer(10, n_via=1, size_mm=5.0, seed=3)
    prob, meshes, m2l = synth.to_reference_objects(board, problem, mesh, geometry)
    layer_solutions = []
    for li in range(len(prob.layers)):
        ms = [m for m, l in zip(meshes, m2l) if l == li]
        forms = []
        for m in ms:
            zf = mesh.ZeroForm(m)
            zf.values[:] = np.random.default_rng(li).random(len(m.vertices))
            forms.append(zf)
        layer_solutions.append(solver.LayerSolution(meshes=ms, potentials=forms, power_densities=[],
                                                    disconnected_meshes=[]))
    solution = solver.Solution(problem=prob, layer_solutions=layer_solutions, solver_info=None)
    paraview.export_solution(solution, tmp_path)
    for li, layer in enumerate(prob.layers):
        root = ET.fromstring((tmp_path / f"{paraview._sanitize_filename(layer.name, set())}.vtu").read_bytes())
        want = []
        for m, zf in zip(layer_solutions[li].meshes, layer_solutions[li].potentials):
            xy = [(vx.p.x, vx.p.y) for vx in m.vertices]
            tri = [[e.origin.i for e in f.edges] for f in m.faces]
            want.append((xy, tri, [zf[vx] for vx in m.vertices]))
        assert _tree(root) == oracle.vtu_file_tree(want)
