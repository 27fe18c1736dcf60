"""ParaView ``.vtu`` export fed from flat arrays -- row F4 of SURVEY.md section 8f.

Mirror of ``padne/paraview.py``: the same ``export_solution(solution, output_dir)`` entry point, the same
file naming (``_sanitize_filename``), the same document -- one ASCII VTK XML UnstructuredGrid file per layer
(``padne/paraview.py:234-298``), one ``<Piece>`` per mesh (``:209-231``).  The reference gathers every
payload by walking vertex and face objects and joins ``str(value)`` strings in Python (``:91, 121-124,
140-145, 159-181``); here the piece text comes from ``pdn_vtu_piece`` (``csrc/pdn_vtu.cpp``) over the arrays
the solve already has -- the coordinates and triangles remembered per mesh (row F2) and the voltage slice
-- so no ``lxml`` tree is built (the reference needs ``lxml``; this module does not).

Bytes: what ``lxml.etree.ElementTree.write(xml_declaration=True, encoding="utf-8", pretty_print=True)``
emits -- the declaration line, two-space indentation, ``<DataArray ...>text</DataArray>`` on one line.  The
element tree (tags, attribute order, payload text) is pinned by ``tests/golden/vtu_*.json.gz``, produced by
the reference's own ``export_solution``; lxml is absent from the image, so its serialiser's whitespace and
declaration spelling are restated from its documented behaviour, not pinned.
"""
from __future__ import annotations

import logging
import re
from pathlib import Path

import numpy as np

from . import capi, flatten

log = logging.getLogger(__name__)

_DECLARATION = b"<?xml version='1.0' encoding='UTF-8'?>\n"
_OPEN = (b'<VTKFile type="UnstructuredGrid" version="0.1" byte_order="LittleEndian">\n'
         b"  <UnstructuredGrid>\n")
_CLOSE = b"  </UnstructuredGrid>\n</VTKFile>\n"


def _sanitize_filename(name: str, used_names: set, fallback_prefix: str = "layer") -> str:
    """File stem for a layer name (``padne/paraview.py:22-59``): characters outside ``[A-Za-z0-9_.-]``
    become ``_``, runs of ``_`` collapse, outer ``_`` are dropped, an empty result is ``fallback_prefix``;
    a stem already in ``used_names`` gets ``_2``, ``_3``, ...  The chosen stem is added to ``used_names``."""
    stem = re.sub(r"_+", "_", re.sub(r"[^a-zA-Z0-9_.-]", "_", (name or "").strip())).strip("_")
    stem = stem or fallback_prefix
    if stem in used_names:
        n = 2
        while f"{stem}_{n}" in used_names:
            n += 1
        stem = f"{stem}_{n}"
    used_names.add(stem)
    return stem


def piece_text(xy, tri, voltage) -> np.ndarray:
    """``create_piece`` (``padne/paraview.py:209-231``) from arrays: uint8 XML text of one ``<Piece>``."""
    return capi.vtu_piece(xy, tri, voltage)


def write_vtu(path, pieces) -> int:
    """One layer file (``padne/paraview.py:270-290``): ``pieces`` is a list of ``(xy, tri, voltage)``.
    Returns the number of bytes written."""
    pieces = list(pieces)
    if not pieces:
        raise ValueError("write_vtu: a layer file needs at least one mesh (the reference skips empty layers)")
    with open(path, "wb") as f:
        n = f.write(_DECLARATION) + f.write(_OPEN)
        for xy, tri, voltage in pieces:
            n += f.write(memoryview(piece_text(xy, tri, voltage)))
        return n + f.write(_CLOSE)


def export_arrays(layer_names, mesh_layer, v_off, t_off, xy, tri, v, output_dir) -> list:
    """The reference's ``export_solution`` on the flat system: meshes back to back as in
    :class:`~padne_b200.flat.FlatSystem` (``v_off`` / ``t_off`` / ``xy`` / ``tri``), ``mesh_layer[m]`` the
    layer of mesh ``m``, ``v`` the solution vector (only ``v[:n_vertices]`` is read).  Layers without a
    mesh are skipped like ``padne/paraview.py:262-264``.  Returns the paths written."""
    output_dir = Path(output_dir)
    output_dir.mkdir(parents=True, exist_ok=True)
    mesh_layer = np.asarray(mesh_layer).reshape(-1)
    v_off = np.asarray(v_off, np.int64)
    t_off = np.asarray(t_off, np.int64)
    xy = np.asarray(xy, np.float64).reshape(-1, 2)
    tri = np.asarray(tri, np.int32).reshape(-1, 3)
    v = np.asarray(v, np.float64).reshape(-1)
    used: set = set()
    written = []
    for li, name in enumerate(layer_names):
        members = np.flatnonzero(mesh_layer == li)
        if members.size == 0:
            log.warning(f"Skipping layer '{name}' - no non-empty meshes")
            continue
        path = output_dir / f"{_sanitize_filename(name, used)}.vtu"
        write_vtu(path, [(xy[v_off[m]:v_off[m + 1]], tri[t_off[m]:t_off[m + 1]], v[v_off[m]:v_off[m + 1]])
                         for m in members])
        written.append(path)
    return written


def export_solution(solution, output_dir: Path) -> None:
    """Drop-in for ``padne.paraview.export_solution`` (``padne/paraview.py:234-298``): one ``.vtu`` per
    layer that has meshes, named after the layer, a piece per mesh with its ``ZeroForm`` values."""
    output_dir = Path(output_dir)
    log.info(f"Exporting solution to ParaView format: {output_dir}")
    output_dir.mkdir(parents=True, exist_ok=True)
    used: set = set()
    total_files = total_pieces = 0
    for layer_idx, layer_solution in enumerate(solution.layer_solutions):
        layer_name = solution.problem.layers[layer_idx].name
        pairs = list(zip(layer_solution.meshes, layer_solution.potentials))
        if not pairs:
            log.warning(f"Skipping layer '{layer_name}' - no non-empty meshes")
            continue
        path = output_dir / f"{_sanitize_filename(layer_name, used)}.vtu"
        pieces = []
        for msh, potential in pairs:
            xy, tri = flatten.mesh_to_arrays(msh)
            pieces.append((xy, tri, potential.values))
        write_vtu(path, pieces)
        total_files += 1
        total_pieces += len(pieces)
    log.info(f"Exported {total_pieces} mesh pieces across {total_files} layer files to {output_dir}")
