"""Golden .vtu element trees from the UNMODIFIED reference exporter (row F4).

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_vtu.py

``padne/paraview.py`` imports ``lxml.etree``, which is absent from this image.  The exporter only uses
``Element`` / ``SubElement`` / ``ElementTree(root).write(path, ...)`` from it, so the stand-in below hands it
``xml.etree.ElementTree``'s element classes and an ``ElementTree`` whose ``write`` records the tree instead of
serialising it.  Everything that decides the CONTENT -- which files, their names, the element structure, the
attribute order and every DataArray payload -- is therefore the reference's own code
(``export_solution`` and the ``create_*`` functions, padne/paraview.py:62-298) run on reference ``Solution``
objects produced by the reference solver.  What lxml's serialiser would add (declaration, indentation) is not
captured.

Stored per case, in ``vtu_<case>.json.gz``: ``{"write_kwargs": {...}, "files": {"<name>.vtu": tree}}`` with
``tree = [tag, [[attr, value], ...], text, [children]]``.  The inputs are the matching
``system_<case>.npz`` (flat arrays and ``v``) plus the layer names stored here.
"""
from __future__ import annotations

import gzip
import json
import os
import sys
import tempfile
import types
import warnings
import xml.etree.ElementTree as ET
from pathlib import Path

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.refload import load_reference  # noqa: E402
from padne_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
problem, mesh, solver = load_reference()
import shapely.geometry as geometry  # noqa: E402  (the loader's stub)

RECORDED = []


class _RecordingTree:
    def __init__(self, root):
        self.root = root

    def write(self, path, **kwargs):
        RECORDED.append((os.path.basename(path), self.root, kwargs))
        Path(path).touch()


def _install_lxml_stand_in():
    lx = types.ModuleType("lxml")
    et = types.ModuleType("lxml.etree")
    et.Element = ET.Element
    et.SubElement = ET.SubElement
    et.ElementTree = _RecordingTree
    lx.etree = et
    sys.modules["lxml"] = lx
    sys.modules["lxml.etree"] = et


def _tree(e):
    return [e.tag, [[k, v] for k, v in e.attrib.items()], e.text, [_tree(c) for c in e]]


def vtu_case(name, board, paraview):
    prob, meshes, m2l = synth.to_reference_objects(board, problem, mesh, geometry)
    vindex = solver.VertexIndexer.create(meshes)
    filtered = list(prob.networks)
    nidx = solver.NodeIndexer.create(prob, meshes, m2l, vindex, filtered)
    L, r = solver.assemble_system(prob, meshes, m2l, vindex, filtered, nidx)
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        v, info = solver.solve_system(L, r)
    layer_solutions = solver.produce_layer_solutions(prob.layers, vindex, meshes, m2l, v,
                                                     [[] for _ in prob.layers])
    solution = solver.Solution(problem=prob, layer_solutions=layer_solutions, solver_info=info)
    del RECORDED[:]
    with tempfile.TemporaryDirectory() as d:
        paraview.export_solution(solution, Path(d))
    kwargs = RECORDED[0][2]
    out = {"layer_names": [l.name for l in prob.layers], "mesh_layer": [int(i) for i in m2l],
           "write_kwargs": kwargs, "files": {fn: _tree(root) for fn, root, _ in RECORDED}}
    with gzip.open(os.path.join(HERE, f"vtu_{name}.json.gz"), "wt", compresslevel=9) as f:
        json.dump(out, f)
    print(f"vtu_{name}: {len(RECORDED)} files, kwargs={kwargs}")


def main():
    _install_lxml_stand_in()
    import importlib
    paraview = importlib.import_module("padne.paraview")
    # same boards and seeds as make_golden.py, so system_<case>.npz holds the matching arrays and v
    vtu_case("via_tht_standin_20", synth.board_via_tht_standin(20, seed=1, n_via=2), paraview)
    vtu_case("4layer_regulator_26", synth.board_4layer_regulator(26, n_via=2, size_mm=12.0, seed=4), paraview)
    vtu_case("8layer_16", synth.board_8layer(16, n_via=1, size_mm=8.0, seed=6), paraview)


if __name__ == "__main__":
    main()
