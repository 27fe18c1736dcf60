"""Row F3 microbench: pdn_nearest_vertex (host buffers, through the C ABI) against the reference's way
(scipy.spatial.KDTree(leafsize=32) per layer built from Python lists + one query per connection,
padne/solver.py:356-396, 421-433) on the 2layer_1M board.

    python profiles/nearest_bench.py [n_queries]
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import scipy.spatial  # noqa: E402

from bench import make_board  # noqa: E402
from padne_b200 import capi  # noqa: E402

n_q = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
fs = make_board("2layer_1M", None).to_flat()
v_off = np.asarray(fs.v_off, np.int64)
xy = np.asarray(fs.xy, np.float64).reshape(-1, 2)
mesh_layer = np.arange(len(v_off) - 1, dtype=np.int32)
rng = np.random.default_rng(0)
q_xy = rng.uniform(xy.min(0), xy.max(0), (n_q, 2))
q_layer = rng.integers(0, len(mesh_layer), n_q).astype(np.int32)
ctx = capi.Context(0)
ctx.nearest_vertex(v_off, mesh_layer, xy, q_xy, q_layer)            # warm-up (buffers)
t0 = time.perf_counter()
reps = 5
for _ in range(reps):
    idx = ctx.nearest_vertex(v_off, mesh_layer, xy, q_xy, q_layer)
t_gpu = (time.perf_counter() - t0) / reps
t0 = time.perf_counter()
ref = np.empty(n_q, np.int64)
for layer in range(len(mesh_layer)):
    pts = [(float(x), float(y)) for x, y in xy[v_off[layer]:v_off[layer + 1]]]      # the reference's list build
    tree = scipy.spatial.KDTree(pts, leafsize=32)
    for q in np.flatnonzero(q_layer == layer):
        _, i = tree.query((q_xy[q, 0], q_xy[q, 1]), k=1)
        ref[q] = v_off[layer] + i
t_cpu = time.perf_counter() - t0
print(json.dumps({"row": "F3 nearest vertex", "vertices": int(xy.shape[0]), "queries": n_q,
                  "gpu_ms_host_buffers": round(1e3 * t_gpu, 3), "cpu_kdtree_ms": round(1e3 * t_cpu, 1),
                  "identical": bool(np.array_equal(idx, ref)),
                  "algorithmic_GBs": round(16.0 * xy.shape[0] / 2 * (n_q / 8) / t_gpu / 1e9, 1)}))
