"""Populate ``oracle/_ref/`` with the reference's own hot-path modules.  TEST / BASELINE INFRASTRUCTURE ONLY.

    python oracle/make_ref.py            (also run by ``__graft_entry__.build()``)

The reference is pure Python, so there is nothing to compile: the recipe copies the three modules of
the hot path -- ``padne/problem.py``, ``padne/mesh.py``, ``padne/solver.py`` --
byte for byte from ``/root/reference`` into ``oracle/_ref/padne/``.  That directory is git-ignored (the
sources never enter this repository's history) but NOT gpurun-ignored, so it travels to the GPU box,
where ``/root/reference`` does not exist.  ``oracle/refload.py`` then imports the modules from there
with the same three stubs as in the build container, and ``bench.py --impl reference`` /
``cpu_baseline`` time the reference's own ``assemble_system`` + ``solve_system``
(``padne/solver.py:783-812, 767-780``) on the box's host cores.

Nothing under ``padne_b200/`` imports ``oracle/_ref``; a manifest with the sha256 of every copied file
is written next to them so a record can say exactly which reference was timed.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("PADNE_REFERENCE_ROOT", "/root/reference")
DEST = os.path.join(HERE, "_ref")
MODULES = ("problem.py", "mesh.py", "solver.py")


def make_ref(verbose: bool = True) -> bool:
    """Copy the modules; returns False (and leaves ``oracle/_ref`` alone) when the reference is absent."""
    src_pkg = os.path.join(REF_SRC, "padne")
    if not all(os.path.isfile(os.path.join(src_pkg, m)) for m in MODULES):
        if verbose:
            print(f"make_ref: {src_pkg} not present; keeping whatever oracle/_ref holds")
        return False
    dst_pkg = os.path.join(DEST, "padne")
    os.makedirs(dst_pkg, exist_ok=True)
    manifest = {"source": REF_SRC, "files": {}}
    for m in MODULES:
        shutil.copyfile(os.path.join(src_pkg, m), os.path.join(dst_pkg, m))
        with open(os.path.join(dst_pkg, m), "rb") as f:
            manifest["files"][f"padne/{m}"] = hashlib.sha256(f.read()).hexdigest()
    sub = os.path.join(REF_SRC, ".SUBMODULES.json")
    if os.path.isfile(sub):
        try:
            manifest["submodules"] = json.load(open(sub))
        except Exception:
            pass
    with open(os.path.join(DEST, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1)
    if verbose:
        print(f"make_ref: copied {', '.join(MODULES)} -> {dst_pkg}")
    return True


if __name__ == "__main__":
    sys.exit(0 if make_ref() else 1)
