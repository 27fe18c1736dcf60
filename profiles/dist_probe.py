"""What does each per-iteration exchange of the partitioned PCG cost?  (torchrun, N ranks)

Runs a fixed number of iterations (rtol unreachable) with none / one / all of the exchanges left out
(PADNE_B200_DIST_SKIP, timing experiment only: the results are garbage) and prints wall-clock per iteration.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from bench import make_board  # noqa: E402
from padne_b200 import capi, dist as pdist  # noqa: E402
from padne_b200.stamps import build_stamps  # noqa: E402

config = sys.argv[1] if len(sys.argv) > 1 else "stack_4layer"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 640
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
fs = make_board(config, None).to_flat()
st = build_stamps(fs)
ctx = capi.Context(local)
pdist.init_comm(ctx)
ctx.assemble(fs.v_off, fs.xy, fs.t_off, fs.tri, fs.conductance, fs.N, *st[:4])


def timed(n_it):
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    try:
        ctx.solve(st[4], n_nodes=fs.n_nodes, rtol=1e-30, max_iter=n_it)
    except capi.NotConvergedError:
        pass
    except Exception as e:  # noqa: BLE001  (a rank that broke down early makes its peers time out)
        return float("nan"), str(e)[:120]
    return time.perf_counter() - t0, None


for skip in ["", "halo", "dots", "rc", "halo,dots,rc"]:
    os.environ["PADNE_B200_DIST_SKIP"] = skip
    timed(32)
    t1, e1 = timed(32)
    t2, e2 = timed(32 + iters)
    if rank == 0:
        print(json.dumps({"config": config, "world": world, "skip": skip or "none", "us_per_iteration": round(1e6 * (t2 - t1) / iters, 1),
                          "error": e1 or e2}), flush=True)
os.environ["PADNE_B200_DIST_SKIP"] = ""
dist.destroy_process_group()
