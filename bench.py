#!/usr/bin/env python
"""bench.py -- DOFs solved/sec (assembly + solve) of the padne FEM hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config NAME] [--n GRID] [--impl reference]

One "step" = one pass of the hot path over one synthetic board: element weights -> CSR assembly ->
constraint elimination -> Krylov solve -> full voltage vector.  At N=1 the workload is
BASELINE.json configs[1] (2-layer 100x100 mm plane, ~1M DOF, one VOLTAGE + one CURRENT source).

  value      whole-job DOF/s with the flat inputs already resident in HBM (device pointers in, device
             pointer out), CUDA events on the library's stream, max over ranks
  e2e        the same metric through the C ABI with HOST buffers: pinned host inputs are copied in,
             the voltage vector is copied out, every step
  roofline   the dominant kernel (SpMV with fused dot products, pdn_kernels.cuh): algorithmic bytes
             per launch / its mean launch duration measured in situ with CUDA events
  cpu_baseline  the oracle port (C assembly restatement + the reference's own scipy spsolve call)
             on a bounded sample of the same board family, 1 core

Multi-GPU (torchrun, one rank per GPU), --multi partitioned (default): ONE board is partitioned across
the ranks -- contiguous row ranges of the reduced system, NCCL halo exchange of the interface entries
before every SpMV and an all-reduce of the fused dot products (csrc/pdn_multi.cu).  Without --config the
board grows with the job, 2 layers of 708x708 (~1M DOF) per GPU, "scaling": "weak"; with an explicit
--config (e.g. 8layer_10M) that fixed board is split, "scaling": "strong".  DOFs are counted once per
step (not once per rank).  --multi independent runs one independent 2layer_1M board per rank instead
(no data-path collective).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "DOFs solved/sec (assembly+solve)"
UNIT = "DOF/s"


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        d = json.load(open(p))
        for k in ("hbm_gbs", "hbm_copy_gbs", "hbm_gb_s"):
            if k in d:
                return float(d[k]), "measured (MEASURED_PEAKS.json)"
        for v in d.values():
            if isinstance(v, dict):
                for k in ("hbm_gbs", "hbm_copy_gbs"):
                    if k in v:
                        return float(v[k]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_board(config: str, n: int | None):
    from padne_b200 import synth
    if config.startswith("stack_"):          # stack_<L>layer: the weak-scaling family, L planes of n x n
        layers = int(config.split("_")[1].replace("layer", ""))
        return synth.board_8layer(n or 708, n_via=16 if (n or 708) >= 200 else 2, size_mm=100.0, n_layers=layers)
    if config == "2layer_1M":
        return synth.board_2layer(n or 708, n_via=16 if (n or 708) >= 200 else 2)
    if config == "4layer_regulator_4M":
        return synth.board_4layer_regulator(n or 1000, n_via=32 if (n or 1000) >= 400 else 2)
    if config == "8layer_10M":
        return synth.board_8layer(n or 1118, n_via=64 if (n or 1118) >= 800 else 2)
    if config == "via_tht_4layer_standin":
        return synth.board_via_tht_standin(n or 64)
    raise SystemExit(f"unknown config {config}")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.rows, self.proc, self.device = [], None, device

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.device}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for nm, val in zip(names, r[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_dofs(config: str, n_sample: int, repeats: int = 1):
    """The oracle port on a bounded sample: C assembly restatement + scipy spsolve (1 core)."""
    from oracle import oracle
    from padne_b200.stamps import build_stamps
    oracle.build()
    fs = make_board(config, n_sample).to_flat()
    stamps = build_stamps(fs)
    best = float("inf")
    t_asm = t_solve = 0.0
    for _ in range(repeats):
        t0 = time.perf_counter()
        indptr, indices, data = oracle.assemble(fs, stamps)
        t1 = time.perf_counter()
        oracle.solve(indptr, indices, data, stamps[4])
        t2 = time.perf_counter()
        if t2 - t0 < best:
            best, t_asm, t_solve = t2 - t0, t1 - t0, t2 - t1
    return fs.N / best, fs.N, t_asm, t_solve


def cpu_reference_dofs(config: str, n_sample: int, repeats: int = 1):
    """The reference's OWN ``assemble_system`` + ``solve_system`` (``padne/solver.py:783-812, 767-780``),
    unmodified, on reference objects built (untimed) by its own ``Mesh.from_triangle_soup`` /
    ``VertexIndexer`` / ``NodeIndexer``.  The modules come from ``/root/reference`` when it is mounted, else
    from ``oracle/_ref`` (``oracle/make_ref.py``).  Returns None when neither is present."""
    from oracle import refload
    if not refload.reference_available():
        return None
    import warnings
    from padne_b200 import synth
    problem, mesh, solver = refload.load_reference()
    import shapely.geometry as geometry          # the loader's stub
    board = make_board(config, n_sample)
    t0 = time.perf_counter()
    prob, meshes, m2l = synth.to_reference_objects(board, problem, mesh, geometry)
    vindex = solver.VertexIndexer.create(meshes)
    nets = list(prob.networks)
    nidx = solver.NodeIndexer.create(prob, meshes, m2l, vindex, nets)
    t_prep = time.perf_counter() - t0
    best, times = None, []
    for _ in range(repeats):
        t0 = time.perf_counter()
        L, r = solver.assemble_system(prob, meshes, m2l, vindex, nets, nidx)
        t1 = time.perf_counter()
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            v, info = solver.solve_system(L, r)
        t2 = time.perf_counter()
        times.append(t2 - t0)
        if best is None or t2 - t0 < best["t"]:
            best = dict(t=t2 - t0, t_asm=t1 - t0, t_solve=t2 - t1, N=int(L.shape[0]),
                        residual_norm=float(info.residual_norm), ground_node_current=float(info.ground_node_current))
    best.update(dofs=best["N"] / best["t"], t_prep=t_prep, origin=refload.reference_origin(), times=times)
    return best


def host_cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU path on the host cores, rank 0 only.  The reference's
    modules (oracle/_ref or /root/reference) are timed when present, else the oracle port."""
    if rank != 0:
        return
    n_sample = args.n or args.ref_sample_n
    vals = []
    ref = cpu_reference_dofs(args.config, n_sample, repeats=args.warmup + args.steps)   # objects built once, untimed
    if ref is not None:
        vals = [(ref["N"] / t, t) for t in ref["times"][args.warmup:]]
        value = float(np.mean([v[0] for v in vals]))
        ms = float(np.mean([v[1] for v in vals]) * 1e3)
        port_dofs, port_N, _, _ = cpu_port_dofs(args.config, n_sample)
        sample = (f"{args.config} board family at grid {n_sample} ({ref['N']} DOF) per step: the reference's own "
                  f"assemble_system ({ref['t_asm']:.2f} s) + solve_system ({ref['t_solve']:.2f} s), unmodified modules from "
                  f"{ref['origin']}; Mesh.from_triangle_soup / VertexIndexer / NodeIndexer ({ref['t_prep']:.2f} s) untimed; "
                  f"1 of {os.cpu_count()} cores ({host_cpu_model()}), the path is serial Python + SuperLU")
        cpu = {"value": value, "unit": UNIT, "cores": 1, "kind": "reference", "sample": sample,
               "residual_norm": ref["residual_norm"], "ground_node_current": ref["ground_node_current"],
               "port_value": port_dofs, "port_note": "oracle C assembly restatement + scipy spsolve on the same sample"}
        N = ref["N"]
    else:
        n_sample = args.n or args.cpu_sample_n
        for k in range(args.warmup + args.steps):
            dofs, N, t_asm, t_solve = cpu_port_dofs(args.config, n_sample)
            if k >= args.warmup:
                vals.append((dofs, t_asm + t_solve))
        value = float(np.mean([v[0] for v in vals]))
        ms = float(np.mean([v[1] for v in vals]) * 1e3)
        sample = (f"{args.config} board family at grid {n_sample} ({N} DOF) per step: oracle C assembly restatement "
                  f"+ the reference's scipy.sparse.linalg.spsolve call (oracle/_ref absent: run oracle/make_ref.py "
                  f"where /root/reference is mounted)")
        cpu = {"value": value, "unit": UNIT, "cores": 1, "kind": "port", "sample": sample}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.config, "grid": n_sample, "dof": N},
            "cpu_baseline": cpu,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def info_dict(info):
    """The solver's own account of one solve, for the JSON line."""
    return {"method": info.method, "preconditioner": info.precond, "iteration_kind": info.iteration_kind,
            "iterations": info.iterations, "restarts": info.restarts, "stagnated": info.stagnated,
            "n_coarse": info.n_coarse, "n_coarse_local": info.n_coarse_local, "local_level": info.local_level,
            "t_coarse_ms": info.t_coarse_ms, "rel_residual": info.rel_residual, "residual_norm": info.residual_norm,
            "ground_node_current": info.ground_node_current, "abs_target": info.abs_target, "floor_abs": info.floor_abs,
            "rhs_norm": info.rhs_norm, "t_krylov_ms": info.t_krylov_ms, "t_setup_ms": info.t_setup_ms,
            "us_per_iteration": 1e3 * info.t_krylov_ms / max(info.iterations, 1),
            "spmv": "x-band staging" if info.spmv_kind == 1 else "tma ring", "spmv_halo_rows": info.spmv_halo,
            "spmv_window_coverage": info.spmv_window_coverage}


class Workload:
    """One synthetic board on this rank: pinned host copies, device copies, and the two timed steps."""

    def __init__(self, torch, capi, ctx, dev, config, n, skw, part=False):
        from padne_b200.stamps import build_stamps
        self.torch, self.ctx, self.config = torch, ctx, config
        fs = make_board(config, n).to_flat()
        s_row, s_col, s_val, s_op, rhs = build_stamps(fs)
        host = dict(v_off=fs.v_off, xy=fs.xy, t_off=fs.t_off, tri=fs.tri, cond=fs.conductance, s_row=s_row,
                    s_col=s_col, s_val=s_val, s_op=s_op, rhs=rhs)
        self.fs = fs
        self.pinned = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in host.items()}
        self.pinned_np = {k: t.numpy() for k, t in self.pinned.items()}
        self.devt = {k: t.to(dev) for k, t in self.pinned.items()}
        self.v_dev = torch.empty(fs.N, dtype=torch.float64, device=dev)
        self.v_host = torch.empty(fs.N, dtype=torch.float64).pin_memory()
        self.v_host_np = self.v_host.numpy()
        self.h2d = int(sum(t.numel() * t.element_size() for t in self.pinned.values()))
        self.d2h = int(fs.N * 8)
        self.skw = dict(skw, n_nodes=fs.n_nodes)
        # several GPUs: partitioned assembly -- every rank builds (and, in the e2e leg, uploads) only its rows' meshes
        self.akw = dict(n_nodes=fs.n_nodes) if part else {}

    def step_device(self):
        d, fs = self.devt, self.fs
        nnz = self.ctx.assemble(d["v_off"], d["xy"], d["t_off"], d["tri"], d["cond"], fs.N, d["s_row"], d["s_col"],
                                d["s_val"], d["s_op"], **self.akw)
        _, info = self.ctx.solve(d["rhs"], out=self.v_dev, **self.skw)
        return nnz, info

    def step_host(self):
        p, fs = self.pinned_np, self.fs
        nnz = self.ctx.assemble(p["v_off"], p["xy"], p["t_off"], p["tri"], p["cond"], fs.N, p["s_row"], p["s_col"],
                                p["s_val"], p["s_op"], **self.akw)
        _, info = self.ctx.solve(p["rhs"], out=self.v_host_np, **self.skw)
        return nnz, info


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default=None, help="default: 2layer_1M at 1 GPU, stack_<2N>layer at N GPUs")
    ap.add_argument("--multi", default="partitioned", choices=["partitioned", "independent"])
    # (--grid: torchrun's own parser trips over "--n" after the script name -- ambiguous with --nnodes ...)
    ap.add_argument("--n", "--grid", dest="n", type=int, default=None,
                    help="grid vertices per side (default: the config's full size)")
    ap.add_argument("--impl", default="padne_b200", choices=["padne_b200", "reference"])
    ap.add_argument("--cpu-sample-n", type=int, default=260, help="grid size of the CPU-baseline sample (oracle port)")
    ap.add_argument("--ref-sample-n", type=int, default=160,
                    help="grid size of the sample the reference's own Python path is timed on (~5 s per step at 160)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-named-configs", action="store_true",
                    help="skip the BASELINE configs 3 and 4 block (4layer_regulator_4M, 8layer_10M)")
    ap.add_argument("--named-steps", type=int, default=3)
    ap.add_argument("--rtol", type=float, default=1e-10)
    ap.add_argument("--atol", type=float, default=0.0, help="absolute bar on ||L v - r||; 0 = the reference's 1e-9, <0 = off")
    ap.add_argument("--precond", default="auto", choices=["auto", "jacobi", "two-level"])
    ap.add_argument("--n-coarse", type=int, default=0)
    ap.add_argument("--replicated-assembly", action="store_true",
                    help="several GPUs: every rank assembles the whole board (round-1 behaviour) instead of its own rows")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "padne_b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    partitioned = world > 1 and args.multi == "partitioned"
    scaling = "weak"
    default_config = args.config is None
    if args.config is None:
        args.config = f"stack_{2 * world}layer" if partitioned else "2layer_1M"
    elif partitioned:
        scaling = "strong"
    args.scaling = scaling
    if args.config.startswith("stack_") and args.n is None:
        # keep the bounded CPU samples at the same DOF count whatever the number of layers
        layers = int(args.config.split("_")[1].replace("layer", ""))
        args.cpu_sample_n = max(40, int(args.cpu_sample_n * (2.0 / layers) ** 0.5))
        args.ref_sample_n = max(30, int(args.ref_sample_n * (2.0 / layers) ** 0.5))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from padne_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ctx = capi.Context(local_rank)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    if partitioned:
        from padne_b200 import dist as pdist
        pdist.init_comm(ctx)                # every pdn_solve on ctx is a collective over all ranks

    precond = {"auto": capi.PRECOND_AUTO, "jacobi": capi.PRECOND_JACOBI, "two-level": capi.PRECOND_TWOLEVEL}[args.precond]
    skw = dict(rtol=args.rtol, atol=args.atol, precond=precond, n_coarse=args.n_coarse)

    def timed(c, fn, steps, warmup, collective):
        """steps timed calls of fn on context c: CUDA events on its stream, L2 flushed before each, max over ranks."""
        stream = ext if c is ctx else torch.cuda.ExternalStream(c.stream, device=dev)
        for _ in range(warmup):
            fn()
        if collective:
            barrier()
        else:
            torch.cuda.synchronize()
        ms, outs = [], []
        launches0 = c.launch_count
        for _ in range(steps):
            flush.zero_()                       # L2 flush between timed iterations (outside the events)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            outs.append(fn())
            e1.record(stream)
            e1.synchronize()
            ms.append(e0.elapsed_time(e1))
        launches = c.launch_count - launches0
        total = torch.tensor([sum(ms)], dtype=torch.float64, device=dev)
        if collective:
            barrier()
            if world > 1:
                dist.all_reduce(total, op=dist.ReduceOp.MAX)
        return float(total.item()), outs, launches

    # ---- the headline workload ---------------------------------------------------------------------
    part_asm = partitioned and not args.replicated_assembly
    wl = Workload(torch, capi, ctx, dev, args.config, args.n, skw, part=part_asm)
    fs = wl.fs
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    t_dev_ms, outs_dev, launches = timed(ctx, wl.step_device, args.steps, args.warmup, True)
    clocks = sampler.stop() if rank == 0 else None
    t_e2e_ms, outs_e2e, _ = timed(ctx, wl.step_host, args.steps, 1, True)
    nnz, info = outs_dev[-1]

    # ---- dominant kernel: SpMV with fused dots (and, round-2 iteration, the restriction), in situ ---------
    n_red = fs.n_nodes
    indptr_tail = ctx.get_csr()[0]
    if part_asm:
        # the resident CSR holds this rank's rows (+ a few terminals): its node entries ARE the launch's share
        nnz_nodes = int(indptr_tail[n_red])
        alg_bytes = 12.0 * nnz_nodes + 20.0 * (n_red / world)
    else:
        nnz_nodes = int(indptr_tail[n_red])
        alg_bytes = 12.0 * nnz_nodes + 20.0 * n_red      # 12 B/nnz + 20 B/row
        if partitioned:
            alg_bytes /= world                           # the timed launch covers this rank's rows only
    del indptr_tail
    spmv_ms = ctx.bench_spmv(reps=200, mode=2)
    spmv_ms_warm = ctx.bench_spmv(reps=200, mode=0)
    spmv_ms_cold = ctx.bench_spmv(reps=50, mode=1)
    peak, peak_src = measured_peak_gbs()
    traffic, traffic_src = None, None
    try:   # dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the committed ncu --set full capture
        t = json.load(open(os.path.join(ROOT, "profiles", "spmv_traffic.json"))).get(args.config)
        if t and world == 1 and args.n is None:
            traffic, traffic_src = t["dram_bytes_read"] + t["dram_bytes_write"], t["source"]
    except Exception:
        pass
    achieved = alg_bytes / (spmv_ms * 1e-3) / 1e9

    # ---- BASELINE configs 3 and 4, inside the same run (put them on the driver's record) ------
    named = {}
    if default_config and not args.no_named_configs and args.n is None:
        del wl.devt, wl.pinned, wl.pinned_np        # ~0.4 GB per GPU back before the 10M board arrives
        torch.cuda.empty_cache()
        # config 4: the 8-layer 300x300 mm backplane, ~10M DOF, strong-scaled over the ranks of this run
        w4 = Workload(torch, capi, ctx, dev, "8layer_10M", None, skw, part=part_asm)
        t4, o4, _ = timed(ctx, w4.step_device, args.named_steps, 2, True)
        i4 = o4[-1][1]
        named["8layer_10M"] = dict(info_dict(i4), dof=w4.fs.N, n_gpus=world, scaling="strong" if world > 1 else "single GPU",
                                   steps=args.named_steps, ms_per_step=t4 / args.named_steps,
                                   value=w4.fs.N * args.named_steps / (t4 * 1e-3), unit=UNIT,
                                   north_star="10M DOF assembled and solved in < 1 s on 8 GPUs")
        del w4
        torch.cuda.empty_cache()
        # config 3: 4 layers, via arrays, two REGULATORs (non-symmetric: BiCGSTAB), ~4M DOF, one GPU (rank 0 alone)
        if rank == 0:
            c3 = capi.Context(local_rank)
            w3 = Workload(torch, capi, c3, dev, "4layer_regulator_4M", None, skw)
            t3, o3, _ = timed(c3, w3.step_device, args.named_steps, 2, False)
            i3 = o3[-1][1]
            named["4layer_regulator_4M"] = dict(info_dict(i3), dof=w3.fs.N, n_gpus=1, steps=args.named_steps,
                                                ms_per_step=t3 / args.named_steps,
                                                value=w3.fs.N * args.named_steps / (t3 * 1e-3), unit=UNIT)
            del w3
            c3.close()
        barrier()

    if world > 1 and os.environ.get("PADNE_B200_PEER_STATS"):
        ctx.comm_destroy()                  # prints the per-exchange wait cycles (collective: every rank)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    dofs_total = fs.N * args.steps * (1 if partitioned else world)
    value = dofs_total / (t_dev_ms * 1e-3)
    e2e_value = dofs_total / (t_e2e_ms * 1e-3)
    kernel_name = ("pdnk::sell_xs_kernel<CG2> (SELL-32-sigma SpMV, x window staged in shared memory, fused r.z, z.w, r.r "
                   "and the restriction P^T w)" if info.iteration_kind == 2 else
                   ("pdnk::sell_xs_kernel<CG>" if info.spmv_kind == 1 else "pdnk::sell_tma_kernel<CG>") +
                   " (SELL-32-sigma SpMV, fused r.z, z.w, r.r)")
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t_dev_ms / args.steps, "higher_is_better": True,
        "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": dict({"workload": args.config, "grid": args.n or "full", "dof": fs.N,
                        "dof_per_gpu": fs.N // world if partitioned else fs.N, "nnz": int(nnz),
                        "layers": fs.n_mesh, "elements": fs.n_elements, "rtol": args.rtol,
                        "atol": args.atol if args.atol != 0 else 1e-9}, **info_dict(info),
                       parallelism=("single GPU" if world == 1 else
                                    f"one board row-partitioned over {world} GPUs (one peer-memory exchange of dots + coarse "
                                    f"sums and one halo exchange per iteration over NVLink, NCCL for setup; "
                                    f"{'assembly and setup partitioned by rows' if part_asm else 'assembly replicated'})"
                                    if partitioned else f"{world} independent boards, one per GPU"),
                       l2="256 MiB written between timed steps; per-step working set exceeds the 126 MB L2"),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h,
                "ms_per_step": t_e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": kernel_name,
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                     "ms_per_launch_in_situ": spmv_ms, "ms_per_launch_warm_l2": spmv_ms_warm,
                     "ms_per_launch_cold_l2": spmv_ms_cold,
                     "note": "algorithmic bytes = SURVEY.md 8d (12 B per stored entry + 20 B per row); in-situ timing "
                             "interleaves the other kernels of a solver iteration, events around the SpMV only"},
    }
    if named:
        line["named_configs"] = named
    if not args.no_cpu_baseline:
        dofs, n_cpu, t_asm, t_solve = cpu_port_dofs(args.config, args.cpu_sample_n)
        port_sample = (f"{args.config} board family at grid {args.cpu_sample_n} ({n_cpu} DOF): oracle C assembly "
                       f"{t_asm:.2f} s + scipy spsolve (the reference's solver call) {t_solve:.2f} s")
        ref = cpu_reference_dofs(args.config, args.ref_sample_n)
        if ref is not None:
            line["cpu_baseline"] = {
                "value": ref["dofs"], "unit": UNIT, "cores": 1, "kind": "reference",
                "sample": f"{args.config} board family at grid {args.ref_sample_n} ({ref['N']} DOF): the reference's own "
                          f"assemble_system {ref['t_asm']:.2f} s + solve_system {ref['t_solve']:.2f} s (unmodified modules, "
                          f"{ref['origin']}); 1 of {os.cpu_count()} cores ({host_cpu_model()}), the path is serial",
                "residual_norm": ref["residual_norm"], "ground_node_current": ref["ground_node_current"],
                "port_value": dofs, "port_sample": port_sample}
        else:
            line["cpu_baseline"] = {"value": dofs, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": port_sample + f"; host has {os.cpu_count()} cores, the path is serial"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
