"""Partitioned solve against the single-GPU solve on the same boards -- needs at least TWO GPUs (skipped otherwise).

One process per GPU under torchrun (``scripts/dist_check.py``): every rank assembles the board alone and then collectively
(partitioned assembly, ``pdn_assemble_part``), solves both ways, and the voltages must agree to 1e-9 relative with
identical bits on all ranks.  Run with and without the rank-local coarse level; with it the iteration
path depends on the partition, so the iteration counts may differ while the voltages may not.
"""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _gpu_count():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("local_level", ["0", "1"])
def test_partitioned_solve_matches_single_gpu(local_level):
    n = _gpu_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    world = 2 if n < 4 else 4
    env = dict(os.environ, PADNE_B200_LOCAL_LEVEL=local_level)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29547", os.path.join(ROOT, "scripts", "dist_check.py"),
           "2layer_1M", "300", "8layer_10M", "300", "4layer_regulator_4M", "120"]
    out = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    rows = [json.loads(line) for line in out.stdout.splitlines() if line.startswith("{")]
    assert len(rows) == 3, out.stdout[-2000:]
    for d in rows:
        assert d["ok"] and d["identical_on_all_ranks"], d
        assert d["max_rel_diff_vs_single"] <= (1e-9 if d["method"] == "pcg" else 1e-7), d   # BiCGSTAB: summation order
        if d["method"] == "pcg":
            assert d["local_level"] == (1 if local_level == "1" else 0), d
            if local_level == "0":
                assert abs(d["iters_multi"] - d["iters_single"]) <= 2, d        # same iteration path without the local level
