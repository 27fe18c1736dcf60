"""Build ``libpadne_b200.so`` in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m padne_b200.build [--force]

The built library lives in ``padne_b200/_lib/`` (git-ignored, NOT gpurun-ignored, so it travels
to the GPU box).  ``pdn_assemble.cu`` is compiled with ``-fmad=false``: the element weights must
round every operation separately like CPython does (``padne/mesh.py:124-139``).
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "_lib")
LIB = os.path.join(LIBDIR, "libpadne_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-I", INCLUDE]
# translation unit -> extra flags
UNITS = {
    "pdn_api.cu": [],
    "pdn_assemble.cu": ["-fmad=false"],
    "pdn_solve.cu": [],
    "pdn_multi.cu": [],
    "pdn_noalloc.cu": [],       # self-contained: pdn_residual_noalloc and its file-scope device scratch
    "pdn_vtu.cpp": [],          # host only (row F4); nvcc hands it to g++
}


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _sources():
    return [u for u in UNITS if os.path.exists(os.path.join(CSRC, u))]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(INCLUDE, "padne_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, defines: list[str] | None = None, tag: str | None = None) -> str:
    """Build the library.  ``defines`` / ``tag`` build a tuning variant ``libpadne_b200_<tag>.so``
    (kernel experiments: select it with the ``PADNE_B200_LIB`` environment variable)."""
    if tag:
        return _build(os.path.join(LIBDIR, f"libpadne_b200_{tag}.so"), verbose, [f"-D{d}" for d in (defines or [])], tag)
    if not force and not needs_build():
        return LIB
    return _build(LIB, verbose, [], "")


def _build(LIB: str, verbose: bool, extra: list[str], tag: str) -> str:
    os.makedirs(LIBDIR, exist_ok=True)
    nvcc = _nvcc()
    objs = []
    procs = []
    for unit in _sources():
        obj = os.path.join(LIBDIR, os.path.splitext(unit)[0] + f"{tag}.o")
        cmd = [nvcc, *ARCH, *COMMON, *UNITS[unit], *extra, "-c", os.path.join(CSRC, unit), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd))
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out)
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed ({p.returncode}): {' '.join(cmd)}\n{out}")
    link = [nvcc, *ARCH, "-shared", "-o", LIB, *objs]
    if verbose:
        print(" ".join(link))
    subprocess.check_call(link)
    for o in objs:
        os.remove(o)
    return LIB


if __name__ == "__main__":
    # python -m padne_b200.build [--force] [-v] [--tag NAME -DFOO=1 ...]
    tag = sys.argv[sys.argv.index("--tag") + 1] if "--tag" in sys.argv else None
    defs = [a[2:] for a in sys.argv if a.startswith("-D")]
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, defines=defs, tag=tag))
