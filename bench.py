#!/usr/bin/env python
"""bench.py -- EVP cell-subcycles/s on the AusCOM 0.1 deg tripole grid (3600 x 2700, ndte = 120).

One "step" = one call of evp(dt): prep kernels, the ndte-subcycle loop, finish kernels.
  value     device-resident: inputs already in HBM, K x evp_gpu_compute, CUDA-event timed
  e2e       the same through the reference-facing call evp_gpu_run (== `call evp(dt)`): pinned
            HOST buffers in, host buffers out, H2D + D2H inside the timed region
  roofline  the fused subcycle kernel against the measured HBM copy bandwidth, 392 algorithmic
            bytes per active cell-subcycle (SURVEY.md section 8d / DESIGN.md)
  cpu_baseline  the CPU restatement (oracle/, `kind: port`) on the host cores, one whole evp call
`--impl reference` times that CPU restatement alone (the reference is Fortran + MPI and cannot be
built in this image; DESIGN.md "Oracle").
Launch: `python bench.py` (N = 1) or `python -m torch.distributed.run --nproc-per-node N bench.py
--gpus N` (one x-slab per GPU, velocity halos over NVLink peer memory, no data-path collective).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

ALG_BYTES = 392.0          # per active cell-subcycle (49 doubles), SURVEY.md section 8d
L2_BYTES = 126e6
# dram__bytes_read.sum + dram__bytes_write.sum of one evp_subcycle_tma_kernel<false> launch on the
# default workload (tp01 full, 1 GPU), from the ncu pass in profiles/r01_launches_tp01_final.md;
# only quoted for exactly that workload
NCU_TRAFFIC = {("tp01", "full", 1, True): 1.7716e9 + 0.8417e9}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--grid", default="tp01")
    ap.add_argument("--scenario", default="full")
    ap.add_argument("--ndte", type=int, default=120)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch the ndte loop kernel by kernel")
    ap.add_argument("--fma", action="store_true", help="use the FMA-contracted comparison build")
    ap.add_argument("--no-compact-metrics", action="store_true",
                    help="do not pass HTE/HTN: fetch all 8 metric planes + tinyarea every subcycle")
    ap.add_argument("--equal-slabs", action="store_true",
                    help="N > 1: nx/N columns per GPU instead of slabs balanced by ocean cells")
    ap.add_argument("--cpu-ndte", type=int, default=120,
                    help="subcycles of the CPU sample (120 = one whole evp call, about 25 s on 16 cores)")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device=0):
        self.p = None
        self.device = device

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


def pinned_like(a):
    import torch
    t = torch.empty(a.shape, dtype=torch.float64 if a.dtype == np.float64 else torch.int32,
                    pin_memory=True)
    v = t.numpy()
    v[...] = a
    return t, v


def cpu_sample(args, case_fn, nthreads=0):
    """Oracle (CPU restatement) on the same workload with fewer subcycles; cell-subcycles/s of the
    ksub loop plus the whole call."""
    from oracle import oracle as orc
    case = case_fn(args.cpu_ndte)
    if nthreads == 0:
        # every core this process may run on, set explicitly: torchrun exports OMP_NUM_THREADS=1
        # to its ranks, which would otherwise leave the reference arm of an N > 1 run on one thread
        nthreads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    t0 = time.perf_counter()
    st = orc.run_case(case, nslabs=1, nthreads=nthreads)
    wall = time.perf_counter() - t0
    cs = st.active_t_cells * args.cpu_ndte
    return {"value": cs / st.total_seconds, "loop_value": cs / st.loop_seconds, "unit": "cell-subcycles/s",
            "seconds": st.total_seconds,
            "cores": nthreads, "kind": "port",
            "sample": f"{args.grid} {args.scenario}, one evp call with ndte={args.cpu_ndte} "
                      f"({cs:.3g} cell-subcycles, {wall:.1f} s wall); oracle/evp_oracle.c, gcc -O2 -fopenmp "
                      f"-ffp-contract=off; the Fortran+MPI reference cannot be built in this image"}


def workload_name(args, dom, dt):
    """config.workload: the same string for both arms."""
    shape = " (AusCOM 0.1 deg shape)" if args.grid == "tp01" else ""
    return f"{args.grid} {dom.nx}x{dom.ny} tripole{shape}, scenario {args.scenario}, ndte={args.ndte}, dt={dt}"


def run_reference(args, rank):
    if rank != 0:
        return
    from cice5_b200 import fixtures as fx
    grid = fx.make_grid(args.grid)
    dt = fx.GRID_DT.get(args.grid, 3600.0)
    vals, loops, secs = [], [], []
    t_all = time.perf_counter()
    for it in range(args.warmup + args.steps):
        if it >= 1 and time.perf_counter() - t_all > 150:      # keep the whole run bounded
            break
        r = cpu_sample(args, lambda n: fx.make_case(grid, args.scenario, ndte=n, dt=dt))
        if it >= min(args.warmup, 1):
            vals.append(r["value"]); loops.append(r["loop_value"]); secs.append(r["seconds"])
    v = statistics.median(vals)
    dom = grid.dom
    out = {"impl": "reference", "metric": "EVP cell-subcycles/s", "value": v, "unit": "cell-subcycles/s",
           "n_gpus": args.gpus, "steps": len(vals), "warmup": min(args.warmup, 1),
           "ms_per_step": 1e3 * statistics.median(secs), "higher_is_better": True, "vs_baseline": None,
           "dtype": "f64", "data": "synthetic",
           "scaling": "strong",
           "config": {"workload": workload_name(args, dom, dt),
                      "decomposition": "one block, OpenMP over the rows (host cores)",
                      "grid_cells": dom.nx * dom.ny,
                      "cpu_sample": ("each step = one whole evp call" if args.cpu_ndte == args.ndte else
                                     f"each step = one evp call with ndte={args.cpu_ndte} instead of {args.ndte}")},
           "cpu_baseline": {**{k: r[k] for k in ("unit", "cores", "kind", "sample")}, "value": v,
                            "loop_value": statistics.median(loops)},
           "e2e": {"value": v, "unit": "cell-subcycles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return 0
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N > 1 with torch.distributed.run --nproc-per-node N")
    import torch
    from cice5_b200 import evp, fixtures as fx

    if evp.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the EVP path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's own banner ("NCCL version ...", printed to
        # stdout whenever the box exports NCCL_DEBUG) goes to stderr instead
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- workload: the named grid split into N x-slabs (the reference's slenderX1 decomposition);
    # total work is fixed as N grows, so scaling is STRONG.
    grid = fx.make_grid(args.grid)
    dom = grid.dom
    dt = fx.GRID_DT.get(args.grid, 3600.0)
    case = fx.make_case(grid, args.scenario, ndte=args.ndte, dt=dt)
    N = world

    # x-slab boundaries: balanced by ocean T cells per column (static, from the land mask -- the
    # reference balances its block distribution the same way), or equal widths
    starts = None
    if N > 1 and not args.equal_slabs:
        starts = fx.balanced_slab_starts(grid.f["tmask"][1:-1, 1:-1] > 0, N)

    def sl(a):
        return fx.slab(a, rank, N, starts) if N > 1 else a

    keep = []

    def pin(d):
        o = {}
        for k, v in d.items():
            t, a = pinned_like(sl(v))
            keep.append(t)
            o[k] = a
        return o

    gridf = {k: sl(grid.f[k]) for k in evp.GRID_R8 + ("tmask", "umask", "HTE", "HTN")}
    inp, io, out = pin(case.inp), pin(case.io), pin(case.out)

    e = evp.EvpGpu.for_domain(dom, args.ndte, dt, rank=rank, nranks=N, device=local_rank,
                              strict=not args.fma, use_cuda_graph=not args.no_graph,
                              compact_metrics=not args.no_compact_metrics, slab_start=starts)
    e.set_grid(gridf)
    if N > 1:
        blob = e.peer_blob()
        blobs = [None] * N
        dist.all_gather_object(blobs, blob)
        e.set_peers(blobs)
    e.upload(inp, io)

    # ---- device-resident: W warm-up + K timed steps ----
    for _ in range(args.warmup):
        e.compute(dt)
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    dev_ms, loop_ms, launches = 0.0, 0.0, 0
    for _ in range(args.steps):
        e.compute(dt)
        st = e.stats()
        dev_ms += st.total_ms
        loop_ms += st.loop_ms
        launches += st.kernel_launches
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    st = e.stats()
    if os.environ.get("EVP_B200_TRACE"):
        # per-subcycle timeline of the last call (device globaltimer): where the time between two
        # subcycle kernels goes at N > 1
        tr = e.trace().astype(np.int64)
        k = tr[:-1]
        med = lambda x: float(np.median(x)) / 1e3
        print(f"[trace rank {rank}] us (median over subcycles; signal..wait only at N > 1): kernel {med(k[:, 1] - k[:, 0]):.1f}  "
              f"signal {med(k[:, 2] - k[:, 1]):.1f}  to-fold {med(k[:, 3] - k[:, 2]):.1f}  "
              f"wait {med(k[:, 4] - k[:, 3]):.1f}  fold+next-launch {med(tr[1:, 0] - k[:, 4]):.1f}  "
              f"period {med(tr[1:, 0] - k[:, 0]):.1f}", file=sys.stderr, flush=True)
    act_t, act_u = st.active_t_cells, st.active_u_cells
    loop_launches = st.loop_kernel_launches

    def allmax(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    dev_ms = allmax(dev_ms)
    loop_ms = allmax(loop_ms)
    wall_ms = allmax(wall_ms)
    # active cells: physical T cells only (the redundant ghost ring is not counted twice)
    act_phys = allsum(float(act_t))
    cs_per_step = act_phys * args.ndte
    value = cs_per_step * args.steps / (dev_ms * 1e-3)

    # ---- e2e through evp_gpu_run with host buffers ----
    e2e = None
    if not args.no_e2e:
        e.evp(dt, inp, io, out)                      # warm
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e.evp(dt, inp, io, out)
        barrier()
        e2e_ms = allmax((time.perf_counter() - t0) * 1e3)
        plane = inp["aice"].nbytes
        iplane = io["iceumask"].nbytes
        # in: 12 fields less ss_tltx/y, which tilt_mode = 0 does not read (not uploaded); inout: u, v, 12 stresses, iceumask
        h2d = 10 * plane + 14 * plane + iplane
        d2h = 15 * plane + iplane + 19 * plane + iplane
        e2e = {"value": cs_per_step * args.steps / (e2e_ms * 1e-3), "unit": "cell-subcycles/s",
               "h2d_bytes_per_step": int(allsum(float(h2d))), "d2h_bytes_per_step": int(allsum(float(d2h))),
               "ms_per_step": e2e_ms / args.steps,
               "note": "evp_gpu_run: pinned host arrays -> H2D -> evp -> D2H of all inout + out fields"}

    if rank == 0:
        peak, peak_src = measured_peaks()
        # the dominant kernel = the fused subcycle kernel; its average duration is the loop's CUDA
        # event time / ndte (the loop also holds the 1-block tripole fold kernel per subcycle)
        sub_ms = loop_ms / (args.steps * args.ndte)
        ach = ALG_BYTES * (act_phys / N) / (sub_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": NCU_TRAFFIC.get((args.grid, args.scenario, N, not args.no_compact_metrics)),
                "traffic_source": "ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, "
                                  "profiles/r01_launches_tp01_final.md (null for other workloads)",
                "peak_source": peak_src,
                "kernel": "evp_subcycle_tma_kernel<false> (fused stress + stepu + E-W halo, TMA ring)"
                          if os.environ.get("EVP_B200_KERNEL", "tma") != "sweep" else "evp_subcycle_kernel<false>",
                "note": "achieved = 392 algorithmic B x active cells of one launch / (ndte-loop CUDA-event time / "
                        "ndte); the loop also holds the small tripole fold kernel of every subcycle.  The kernel "
                        "really moves fewer bytes than the algorithmic 392 B/cell (compact metric rows, merged "
                        "coefficient planes): see traffic",
                "alg_bytes_per_cell_subcycle": ALG_BYTES, "avg_launch_ms": sub_ms,
                "active_cells_per_launch": act_phys / N}
        res = {"metric": "EVP cell-subcycles/s", "value": value, "unit": "cell-subcycles/s",
               "n_gpus": N, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
               "config": {"workload": workload_name(args, dom, dt),
                          "decomposition": (f"{N} x-slab(s) of {dom.nx // N} columns" if starts is None else
                                            f"{N} x-slabs balanced by ocean cells, first columns {starts}"),
                          "active_t_cells": act_phys, "grid_cells": dom.nx * dom.ny,
                          "l2": "working set per subcycle (~3 GB) exceeds the 126 MB L2; no flush needed"},
               "loop_ms_per_step": loop_ms / args.steps, "wall_ms_per_step": wall_ms / args.steps,
               "grid_value": dom.nx * dom.ny * args.ndte * args.steps / (dev_ms * 1e-3),
               "roofline": roof, "clocks": clocks, "e2e": e2e,
               "gpu_launches": int(launches), "loop_kernel_launches_per_step": int(loop_launches)}
        if not args.no_cpu_baseline and N == 1:
            try:
                res["cpu_baseline"] = cpu_sample(
                    args, lambda n: fx.make_case(grid, args.scenario, ndte=n, dt=dt))
            except Exception as ex:       # the oracle is only the reported baseline
                res["cpu_baseline"] = {"value": None, "error": str(ex)}
        print(json.dumps(res), flush=True)
    e.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
