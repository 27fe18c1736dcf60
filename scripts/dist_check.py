"""Multi-GPU parity check (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/dist_check.py [config grid]...

Every rank assembles the same board, solves it alone (single-GPU context) and then collectively
(row-partitioned Krylov loop, NCCL halo exchange + all-reduce); the two voltage vectors must agree
to 1e-9 relative and all ranks must return identical bits.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from bench import make_board  # noqa: E402
from padne_b200 import capi, dist as pdist  # noqa: E402
from padne_b200.stamps import build_stamps  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cases = sys.argv[1:] or ["2layer_1M", "120", "4layer_regulator_4M", "90", "8layer_10M", "200"]
single = capi.Context(local)
multi = capi.Context(local)
pdist.init_comm(multi)
ok = True
for name, grid in zip(cases[0::2], cases[1::2]):
    fs = make_board(name, int(grid)).to_flat()
    st = build_stamps(fs)
    single.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])
    part = os.environ.get("DIST_CHECK_REPLICATED", "0") != "1"       # default: partitioned assembly (pdn_assemble_part)
    multi.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4], **(dict(n_nodes=fs.n_nodes) if part else {}))
    v1, i1 = single.solve(st[4], n_nodes=fs.n_nodes)
    dist.barrier()
    v2, i2 = multi.solve(st[4], n_nodes=fs.n_nodes)
    scale = np.max(np.abs(v1[:fs.n_nodes]))
    err = float(np.max(np.abs(v1[:fs.n_nodes] - v2[:fs.n_nodes])) / scale)
    # identical bits on every rank
    t = torch.from_numpy(v2).cuda()
    t0 = t.clone()
    dist.broadcast(t0, src=0)
    same = bool(torch.equal(t, t0))
    flag = torch.tensor([int(same)], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    # PCG takes the identical iteration path on any number of ranks (1e-9); BiCGSTAB is
    # sensitive to the summation order of its dot products, so two converged runs only agree to the
    # accuracy the 1e-10 residual buys (measured 8e-9)
    # With the rank-local coarse level (default on several GPUs) the preconditioner, hence the iteration path,
    # depends on the partition: PADNE_B200_LOCAL_LEVEL=0 restores the identical path and the 1e-9 bound.
    same_path = (i2.method == "pcg" and i2.iteration_kind == i1.iteration_kind and
                 os.environ.get("PADNE_B200_LOCAL_LEVEL", "1" if world >= 4 else "0")[:1] == "0")
    tol = 1e-9 if same_path else 1e-7
    good = err < tol and bool(flag.item()) and i2.rel_residual <= 1.05e-10
    ok = ok and good
    if rank == 0:
        print(json.dumps({"case": name, "grid": int(grid), "N": fs.N, "world": world, "method": i2.method,
                          "partitioned_assembly": part, "residual_norm_single": i1.residual_norm,
                          "iters_single": i1.iterations, "iters_multi": i2.iterations, "kind_multi": i2.iteration_kind,
                          "n_coarse": i2.n_coarse, "n_coarse_local": i2.n_coarse_local, "local_level": i2.local_level,
                          "us_per_iter_multi": round(1e3 * i2.t_krylov_ms / max(i2.iterations, 1), 1),
                          "krylov_ms_single": round(i1.t_krylov_ms, 2), "krylov_ms_multi": round(i2.t_krylov_ms, 2),
                          "rel_residual_multi": i2.rel_residual, "residual_norm_multi": i2.residual_norm,
                          "max_rel_diff_vs_single": err, "identical_on_all_ranks": bool(flag.item()), "ok": good}),
              flush=True)
dist.barrier()
multi.comm_destroy()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
