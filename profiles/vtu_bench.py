"""Row F4 timing (CPU, build container): padne_b200.paraview.export_solution against the reference's own
export_solution (padne/paraview.py:234-298) on the same reference Solution objects.  lxml is absent here, so
the reference runs with the recording stand-in of tests/golden/make_golden_vtu.py: its time covers the
object walks and string joins but NOT lxml's serialisation and file write, i.e. it is a lower bound.

    python profiles/vtu_bench.py [grid_side]      # default 224 -> ~100k DOF on two layers
"""
import os
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import make_golden_vtu as g  # noqa: E402  (loads the reference, installs nothing yet)
from padne_b200 import paraview as ours, synth  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 224
g._install_lxml_stand_in()
import importlib  # noqa: E402
ref_paraview = importlib.import_module("padne.paraview")

board = synth.board_2layer(side, n_via=4, size_mm=100.0, seed=0)
prob, meshes, m2l = synth.to_reference_objects(board, g.problem, g.mesh, g.geometry)
rng = np.random.default_rng(0)
layer_solutions = []
for li in range(len(prob.layers)):
    ms = [m for m, l in zip(meshes, m2l) if l == li]
    forms = []
    for m in ms:
        zf = g.mesh.ZeroForm(m)
        zf.values[:] = rng.random(len(m.vertices))
        forms.append(zf)
    layer_solutions.append(g.solver.LayerSolution(meshes=ms, potentials=forms, power_densities=[],
                                                  disconnected_meshes=[]))
solution = g.solver.Solution(problem=prob, layer_solutions=layer_solutions, solver_info=None)
n_v = sum(len(m.vertices) for m in meshes)

with tempfile.TemporaryDirectory() as d:
    t = time.perf_counter()
    ref_paraview.export_solution(solution, Path(d))
    t_ref = time.perf_counter() - t
with tempfile.TemporaryDirectory() as d:
    t = time.perf_counter()
    ours.export_solution(solution, Path(d))          # first call walks the object graph once (row F2 caches it)
    t_cold = time.perf_counter() - t
    t = time.perf_counter()
    ours.export_solution(solution, Path(d))
    t_warm = time.perf_counter() - t
    size = sum(os.path.getsize(os.path.join(d, f)) for f in os.listdir(d))
print(f"vertices {n_v}  reference (no serialise/write) {t_ref:.3f} s  "
      f"ours cold (flatten walk + write) {t_cold:.3f} s  ours with remembered arrays {t_warm:.3f} s  "
      f"{size / 1e6:.1f} MB written")
