"""Load the UNMODIFIED reference padne modules (problem, mesh, solver) for oracle work.

TEST INFRASTRUCTURE ONLY.  Nothing in ``padne_b200/`` may import this module.

The reference lives at ``/root/reference`` in the build container.  It is not present on the GPU box;
there the loader falls back to ``oracle/_ref/`` -- the byte-for-byte copy of the three hot-path modules
made by ``oracle/make_ref.py`` (git-ignored, shipped with the snapshot).  Users:

* ``tests/golden/make_golden*.py`` -- generate the committed golden fixtures (build container);
* ``tests/test_dropin.py``, ``tests/test_halfedge.py``, ``tests/test_oracle.py`` -- pin the oracle and the
  host mirror against the real reference (skipped automatically when no reference is found);
* ``bench.py --impl reference`` / ``cpu_baseline`` -- time the reference's own ``assemble_system`` +
  ``solve_system`` (``padne/solver.py:783-812, 767-780``) on the host cores (``kind: "reference"``).

Three stubs make ``padne/problem.py``, ``padne/mesh.py`` and ``padne/solver.py`` import
without KiCad, shapely or the CGAL extension:

1. a bare ``padne`` package object (bypasses ``padne/__init__.py:2`` which imports kicad);
2. a stub ``padne._cgal`` (``mesh.py:4``);
3. a stub ``shapely`` exposing the few names the three modules touch at import time and
   ``shapely.geometry.Point(x, y)`` with ``.x`` / ``.y`` (``problem.py:51``, ``solver.py:425``).
"""
from __future__ import annotations

import importlib
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))


def _find_root() -> str:
    for cand in (os.environ.get("PADNE_REFERENCE_ROOT"), "/root/reference", os.path.join(_HERE, "_ref")):
        if cand and os.path.isfile(os.path.join(cand, "padne", "solver.py")):
            return cand
    return os.environ.get("PADNE_REFERENCE_ROOT", "/root/reference")


REFERENCE_ROOT = _find_root()


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "padne", "solver.py"))


def reference_origin() -> str:
    """Where the loaded modules come from: the mounted reference or the shipped copy of its files."""
    return "oracle/_ref (copy made by oracle/make_ref.py)" if REFERENCE_ROOT.startswith(_HERE) else REFERENCE_ROOT


class _StubPoint:
    def __init__(self, x, y):
        self.x = float(x)
        self.y = float(y)


class _StubMultiPolygon:
    def __init__(self, geoms=()):
        self.geoms = tuple(geoms)


class _StubPolygon:
    def __init__(self, *a, **k):
        pass


def _install_stubs() -> None:
    if "shapely" not in sys.modules:
        sh = types.ModuleType("shapely")
        geom = types.ModuleType("shapely.geometry")
        strtree = types.ModuleType("shapely.strtree")
        geom.Point = _StubPoint
        geom.Polygon = _StubPolygon
        geom.MultiPolygon = _StubMultiPolygon
        geom.LinearRing = _StubPolygon
        strtree.STRtree = type("STRtree", (), {})
        sh.geometry = geom
        sh.strtree = strtree
        sh.Point = _StubPoint
        sh.__padne_oracle_stub__ = True
        sys.modules["shapely"] = sh
        sys.modules["shapely.geometry"] = geom
        sys.modules["shapely.strtree"] = strtree
    if "padne" not in sys.modules:
        pkg = types.ModuleType("padne")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "padne")]
        sys.modules["padne"] = pkg
    if "padne._cgal" not in sys.modules:
        cg = types.ModuleType("padne._cgal")
        cg.PolyBoundaryDistanceMap = type("PolyBoundaryDistanceMap", (), {})
        cg.CGALPolygon = type("CGALPolygon", (), {})
        sys.modules["padne._cgal"] = cg


def load_reference():
    """Return ``(problem, mesh, solver)`` -- the reference's own modules, unmodified."""
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    _install_stubs()
    problem = importlib.import_module("padne.problem")
    mesh = importlib.import_module("padne.mesh")
    solver = importlib.import_module("padne.solver")
    return problem, mesh, solver
