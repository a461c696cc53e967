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
# dram__bytes_read.sum + dram__bytes_write.sum PER SUBCYCLE of evp_loop_tma_kernel on the default workload
# (tp01 full, 1 GPU: 219.20 + 104.83 GB over the 119 subcycles of one launch), from the ncu --set full capture in
# profiles/r02_loop_ncu.md; only quoted for exactly that workload
NCU_TRAFFIC = {("tp01", "full", 1, True): (219.202653e9 + 104.833455e9) / 119.0}


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
    ap.add_argument("--no-parity", action="store_true", help="skip the bitwise comparison with the oracle")
    ap.add_argument("--no-transport", action="store_true",
                    help="skip the transport-stage block (departure_points + field reconstruction, N = 1 only)")
    ap.add_argument("--no-other-configs", action="store_true",
                    help="skip the short measurements of the other BASELINE grid shapes")
    ap.add_argument("--no-eap", action="store_true", help="skip the EAP (kdyn = 2) block (N = 1 only)")
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
    """nvidia-smi in a loop beside the run.  It is started before the warm-up (a query takes tens of ms and a timed
    region at N = 8 is only ~0.1-0.2 s long) and the samples are selected by their time stamps: those inside the
    timed region, or -- when fewer than two fall inside it -- those of warm-up + timed region (same kernels, same load)."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device=0):
        self.p = None
        self.device = device
        self.t0 = self.t1 = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "25"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def mark_begin(self):
        import datetime
        self.t0 = datetime.datetime.now()

    def mark_end(self):
        import datetime
        self.t1 = datetime.datetime.now()

    def stop(self):
        import datetime
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        rows = []
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f")
                rows.append((ts, float(f[1]), float(f[2]), [n for n, v in zip(names, f[4:8]) if v.lower().startswith("active")]))
            except ValueError:
                continue
        inside = [r for r in rows if self.t0 and self.t1 and self.t0 <= r[0] <= self.t1]
        window = "timed region"
        if len(inside) < 2:
            inside = [r for r in rows if not self.t1 or r[0] <= self.t1]
            window = "warm-up + timed region (the timed region is shorter than two nvidia-smi queries)"
        sm = [r[1] for r in inside]
        reasons = sorted({n for r in inside for n in r[3]})
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(r[2] for r in inside) if inside else None, "samples": len(sm),
                "window": window, "reasons": reasons}


def pinned_like(a):
    import torch
    t = torch.empty(a.shape, dtype=torch.float64 if a.dtype == np.float64 else torch.int32,
                    pin_memory=torch.cuda.is_available())
    v = t.numpy()
    v[...] = a
    return t, v


def cpu_sample(args, case_fn, nthreads=0, keep=None):
    """Oracle (CPU restatement) on the same workload with fewer subcycles; cell-subcycles/s of the
    ksub loop plus the whole call.  keep: a list that receives the finished case (the parity leg reuses it)."""
    from oracle import oracle as orc
    case = case_fn(args.cpu_ndte)
    if keep is not None:
        keep.append(case)
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


SHAPES = {"gx3": "displaced pole, real grid + kmt (BASELINE configs[0])", "gx1": "displaced pole, real kmt (configs[1])",
          "tp1": "tripole, ACCESS-OM 1 deg shape (configs[2])", "tp025": "tripole, ACCESS-OM 0.25 deg shape (configs[3])",
          "tp01": "tripole, AusCOM 0.1 deg shape (configs[4])", "tp01s8": "tripole, one eighth of tp01 (tuning proxy)"}


def workload_name(args, dom, dt, grid=None):
    """config.workload: the same string for both arms."""
    g = grid or args.grid
    return f"{g} {dom.nx}x{dom.ny} {SHAPES.get(g, '')}, scenario {args.scenario}, ndte={args.ndte}, dt={dt}"


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


PARITY_IO = ("uvel", "vvel") + tuple(f"stress{k}_{c}" for k in ("p", "m", "12") for c in (1, 2, 3, 4)) + ("fm", "iceumask")
PARITY_OUT = ("strairx", "strairy", "strtltx", "strtlty", "strocnx", "strocny", "strintx", "strinty", "strocnxT",
              "strocnyT", "divu", "shear", "rdg_conv", "rdg_shear", "prs_sig", "uvel_init", "vvel_init", "tmass",
              "strength_out", "icetmask")


def _digest(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a[1:-1, 1:-1]).tobytes()).hexdigest()


class Ctx:
    """torch.distributed plumbing of one bench process (control plane only: the data plane is peer memory)."""

    def __init__(self):
        import torch
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        torch.cuda.set_device(self.local_rank)
        self.dist = None
        if self.world > 1:
            # stdout carries exactly one JSON line: NCCL's own banner goes to stderr
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            import torch.distributed as dist_
            self.dist = dist_
            dist_.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def _red(self, x, op):
        if self.dist is None:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def allmax(self, x):
        return self._red(x, self.dist.ReduceOp.MAX if self.dist else None)

    def allmin(self, x):
        return self._red(x, self.dist.ReduceOp.MIN if self.dist else None)

    def allsum(self, x):
        return self._red(x, self.dist.ReduceOp.SUM if self.dist else None)

    def gather(self, obj):
        if self.dist is None:
            return [obj]
        out = [None] * self.world if self.rank == 0 else None
        self.dist.gather_object(obj, out, dst=0)
        return out

    def bcast(self, obj):
        if self.dist is None:
            return obj
        box = [obj]
        self.dist.broadcast_object_list(box, src=0)
        return box[0]


def measure(ctx, args, gridname, scenario, steps, warmup, do_e2e, do_parity, ref_case=None, traffic=None):
    """One workload on the N ranks of this run: device-resident throughput, optional e2e through evp_gpu_run,
    optional bitwise comparison of one whole evp call (ndte as benchmarked, same handles / peers / graph /
    streams as the timed legs) with the oracle.  Returns the result dictionary on rank 0 (None elsewhere)."""
    from cice5_b200 import evp, fixtures as fx
    rank, N, local_rank = ctx.rank, ctx.world, ctx.local_rank
    grid = fx.make_grid(gridname)
    dom = grid.dom
    dt = fx.GRID_DT.get(gridname, 3600.0)
    case = fx.make_case(grid, scenario, ndte=args.ndte, dt=dt)

    # x-slab boundaries: balanced by ocean T cells per column (static, from the land mask -- the
    # reference balances its block distribution the same way), or equal widths
    starts = None
    if N > 1 and (not args.equal_slabs or dom.nx % N):
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
    io0 = {k: v.copy() for k, v in io.items()}          # the state before the first call

    e = evp.EvpGpu.for_domain(dom, args.ndte, dt, rank=rank, nranks=N, device=local_rank,
                              strict=not args.fma, use_cuda_graph=not args.no_graph,
                              compact_metrics=not args.no_compact_metrics, slab_start=starts)
    e.set_grid(gridf)
    if N > 1:
        blobs = [None] * N
        ctx.dist.all_gather_object(blobs, e.peer_blob())
        e.set_peers(blobs)

    # ---- parity: ONE whole call from the initial state, every inout / out field, bitwise ----
    parity = None
    if do_parity:
        e.upload(inp, io0)
        e.compute(dt)
        pio = {k: np.zeros_like(v) for k, v in io.items()}
        pout = {k: np.zeros_like(v) for k, v in out.items()}
        e.download(pio, pout)
        mine = {k: _digest(pio[k]) for k in PARITY_IO}
        mine.update({k: _digest(pout[k]) for k in PARITY_OUT})
        umax_p = float(max(np.abs(pio["uvel"]).max(), np.abs(pio["vvel"]).max()))
        allh = ctx.gather(mine)
        verdict = None
        if rank == 0:
            from oracle import oracle as orc
            t0 = time.perf_counter()
            if ref_case is None:
                ref_case = fx.make_case(grid, scenario, ndte=args.ndte, dt=dt)
                orc.run_case(ref_case, nslabs=1, nthreads=len(os.sched_getaffinity(0)))
            bad = []
            for r in range(N):
                for k in PARITY_IO + PARITY_OUT:
                    ref = (ref_case.io if k in PARITY_IO else ref_case.out)[k]
                    ref = fx.slab(ref, r, N, starts) if N > 1 else ref
                    if _digest(ref) != allh[r][k]:
                        bad.append((r, k))
            verdict = {"bitwise": not bad, "fields": len(PARITY_IO) + len(PARITY_OUT), "nranks": N, "ndte": args.ndte,
                       "mismatches": bad[:20], "oracle_s": round(time.perf_counter() - t0, 1),
                       "what": "one whole evp call from the initial state through the benchmarked handles, every "
                               "inout/out field of every rank's slab, SHA-256 of the physical cells against "
                               "oracle/evp_oracle.c run as one block"}
        verdict = ctx.bcast(verdict)
        max_rel = 0.0
        if not verdict["bitwise"]:
            # diagnosis only: the relative size of the differences, field by field on the failing ranks
            refs = ctx.bcast({k: (ref_case.io if k in PARITY_IO else ref_case.out)[k] for k in PARITY_IO + PARITY_OUT}
                             if rank == 0 else None)
            for (r, k) in verdict["mismatches"]:
                if r == rank:
                    ref = sl(refs[k])[1:-1, 1:-1].astype("f8")
                    got = (pio if k in PARITY_IO else pout)[k][1:-1, 1:-1].astype("f8")
                    den = np.abs(ref).max()
                    max_rel = max(max_rel, float(np.abs(got - ref).max() / den) if den > 0 else float("inf"))
        verdict["max_rel"] = ctx.allmax(max_rel)
        verdict["umax"] = ctx.allmax(umax_p)
        parity = verdict

    # ---- device-resident: W warm-up + K timed steps ----
    e.upload(inp, io0)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(warmup):
        e.compute(dt)
    ctx.barrier()
    sampler.mark_begin()
    t0 = time.perf_counter()
    dev_ms, loop_ms, launches = 0.0, 0.0, 0
    for _ in range(steps):
        e.compute(dt)
        st = e.stats()
        dev_ms += st.total_ms
        loop_ms += st.loop_ms
        launches += st.kernel_launches
    ctx.barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else None
    st = e.stats()
    if os.environ.get("EVP_B200_TRACE") and st.loop_kernel_launches > 8:
        tr = e.trace().astype(np.int64)
        k = tr[:-1]
        med = lambda x: float(np.median(x)) / 1e3
        print(f"[trace rank {rank}] us (median over subcycles; signal..wait only at N > 1): kernel {med(k[:, 1] - k[:, 0]):.1f}  "
              f"signal {med(k[:, 2] - k[:, 1]):.1f}  to-fold {med(k[:, 3] - k[:, 2]):.1f}  "
              f"wait {med(k[:, 4] - k[:, 3]):.1f}  fold+next-launch {med(tr[1:, 0] - k[:, 4]):.1f}  "
              f"period {med(tr[1:, 0] - k[:, 0]):.1f}", file=sys.stderr, flush=True)
    act_t = st.active_t_cells
    loop_launches = st.loop_kernel_launches
    # the state after W + K calls: finite, and moving
    sio = {k: np.zeros_like(v) for k, v in io.items()}
    e.download(sio, None)
    fin_v = all(bool(np.isfinite(sio[k]).all()) for k in PARITY_IO if k != "iceumask")
    umax_v = float(max(np.abs(sio["uvel"]).max(), np.abs(sio["vvel"]).max()))

    dev_ms_r, loop_ms_r = dev_ms, loop_ms
    dev_ms = ctx.allmax(dev_ms)
    loop_ms = ctx.allmax(loop_ms)
    loop_ms_min = ctx.allmin(loop_ms_r)
    wall_ms = ctx.allmax(wall_ms)
    act_max = ctx.allmax(float(act_t))
    act_phys = ctx.allsum(float(act_t))        # physical T cells only (the redundant ghost ring is not counted twice)
    cs_per_step = act_phys * args.ndte
    value = cs_per_step * steps / (dev_ms * 1e-3)
    check_value = {"all_finite": ctx.allmin(1.0 if fin_v else 0.0) == 1.0, "umax": ctx.allmax(umax_v)}

    # ---- e2e through evp_gpu_run (== `call evp(dt)`) with HOST buffers, copies inside the timed region ----
    #   e2e       resident-state protocol (what the Fortran shim does by default): per call the 10 input arrays
    #             go up (aice, vice, vsno, aice_init, strairxT/yT, uocn, vocn, Cdn_ocn, strength), the dynamic
    #             state stays on the device, and what the next host phases read comes back (uvel, vvel -> transport;
    #             divu, shear, rdg_conv, rdg_shear -> ridging; strocnxT/yT -> coupler): source/ice_step_mod.F90:1126-1141
    #   e2e_full  every inout and out array up and down on every call (36 arrays back), as in round 1
    e2e = e2e_full = None
    if do_e2e:
        plane = inp["aice"].nbytes
        iplane = io["iceumask"].nbytes

        def timed(resident):
            for k, v in io0.items():
                io[k][...] = v
            e.invalidate_state()
            e.set_resident(resident, (evp.DL_VELOCITY | evp.DL_DEFORM | evp.DL_OCN_T) if resident else evp.DL_ALL)
            e.evp(dt, inp, io, out)                      # warm (resident: the one call that uploads the state)
            ctx.barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                e.evp(dt, inp, io, out)
            ctx.barrier()
            ms = ctx.allmax((time.perf_counter() - t0) * 1e3)
            names_io = ("uvel", "vvel") if resident else tuple(k for k in PARITY_IO if k != "iceumask")
            names_out = ("divu", "shear", "rdg_conv", "rdg_shear", "strocnxT", "strocnyT") if resident else \
                tuple(k for k in PARITY_OUT if k != "icetmask")
            fin = all(bool(np.isfinite(io[k]).all()) for k in names_io) and \
                all(bool(np.isfinite(out[k]).all()) for k in names_out)
            umax = float(max(np.abs(io["uvel"]).max(), np.abs(io["vvel"]).max()))
            if resident:
                h2d, d2h = 10 * plane, 8 * plane
                note = ("evp_gpu_run, resident-state protocol: pinned host arrays -> H2D of the 10 per-call inputs -> evp "
                        "-> D2H of uvel, vvel, divu, shear, rdg_conv, rdg_shear, strocnxT, strocnyT")
            else:
                # in: 12 fields less ss_tltx/y (tilt_mode = 0 does not read them); inout: u, v, 12 stresses, iceumask
                h2d = 10 * plane + 14 * plane + iplane
                d2h = 15 * plane + iplane + 19 * plane + iplane
                note = "evp_gpu_run, full copy: H2D of all in + inout arrays -> evp -> D2H of all inout + out arrays"
            return {"value": cs_per_step * steps / (ms * 1e-3), "unit": "cell-subcycles/s",
                    "h2d_bytes_per_step": int(ctx.allsum(float(h2d))), "d2h_bytes_per_step": int(ctx.allsum(float(d2h))),
                    "ms_per_step": ms / steps, "all_finite": ctx.allmin(1.0 if fin else 0.0) == 1.0,
                    "umax": ctx.allmax(umax), "note": note}

        e2e_full = timed(False)
        e2e = timed(True)
        e.set_resident(False, evp.DL_ALL)

    res = None
    if rank == 0:
        peak, peak_src = measured_peaks()
        # the dominant kernel: the persistent loop kernel runs all ndte subcycles in one launch, so "one launch
        # of the old per-subcycle kernel" is the loop's CUDA-event time / ndte
        sub_ms = loop_ms / (steps * args.ndte)
        ach = ALG_BYTES * act_max / (sub_ms * 1e-3) / 1e9
        loop_kernel = loop_launches <= 8        # loop kernel + last subcycle + fold (+ signal / wait pairs at N > 1)
        roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "peak_source": peak_src,
                "kernel": ("evp_loop_tma_kernel (subcycles 1 .. ndte-1 in one launch: fused stress + stepu + halo, TMA ring) + evp_subcycle_tma_kernel<true> (the last one)" if loop_kernel
                           else "evp_subcycle_tma_kernel<false>" if os.environ.get("EVP_B200_KERNEL", "tma") != "sweep"
                           else "evp_subcycle_kernel<false>"),
                "note": "achieved = 392 algorithmic B x active cells of the busiest rank x ndte / CUDA-event time of the "
                        "ndte loop; traffic = ncu DRAM bytes per subcycle of the same workload (profiles/), null if not "
                        "captured for it",
                "alg_bytes_per_cell_subcycle": ALG_BYTES, "avg_subcycle_ms": sub_ms,
                "active_cells_per_launch": act_max}
        res = {"metric": "EVP cell-subcycles/s", "value": value, "unit": "cell-subcycles/s",
               "n_gpus": N, "steps": steps, "warmup": warmup,
               "ms_per_step": dev_ms / steps, "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
               "config": {"workload": workload_name(args, dom, dt, gridname),
                          "decomposition": (f"{N} x-slab(s) of {dom.nx // N} columns" if starts is None else
                                            f"{N} x-slabs balanced by ocean cells, first columns {starts}"),
                          "active_t_cells": act_phys, "grid_cells": dom.nx * dom.ny,
                          "l2": ("working set per subcycle exceeds the 126 MB L2; no flush needed"
                                 if ALG_BYTES * act_max > L2_BYTES else
                                 "working set per subcycle fits the L2 (an L2-resident configuration by nature: the "
                                 "same state is swept ndte times per call); reported for completeness")},
               "loop_ms_per_step": loop_ms / steps, "loop_ms_per_step_fastest_rank": loop_ms_min / steps,
               "wall_ms_per_step": wall_ms / steps,
               "us_per_subcycle": 1e3 * sub_ms,
               "grid_value": dom.nx * dom.ny * args.ndte * steps / (dev_ms * 1e-3),
               "roofline": roof, "clocks": clocks, "value_check": check_value, "parity": parity, "e2e": e2e,
               "e2e_full": e2e_full,
               "gpu_launches": int(launches), "loop_kernel_launches_per_step": int(loop_launches)}
    # the bench workload: timing only (its oracle twin would hold 370 planes of 78 MB on the host); the 1-degree tripole
    # among the other configurations: timing + the oracle port beside it (parity of every cell, CPU baseline)
    if N == 1 and (do_e2e or gridname == "tp1") and not getattr(args, "no_transport", False) and res is not None:
        try:
            res["transport"] = transport_block(e, args, grid, case, dt, cpu=not do_e2e)
        except Exception as ex:                   # noqa: BLE001 -- a side block must not cost the headline line
            res["transport"] = {"error": str(ex)[:300]}
    e.close()
    return res


def transport_block(e, args, grid, case, dt, steps=5, warmup=2, cpu=False):
    """SURVEY.md 8f row 2, the stages of horizontal_remap that run on the device so far, on the bench workload's grid
    (N = 1): departure_points from the velocities the timed evp calls left resident, and the field reconstruction
    (make_masks + construct_fields + limited_gradient) of 5 ice categories x 11 tracers + open water.  Device time by
    CUDA events inside the library (evp_gpu_get_transport_timing); bandwidth against the same measured HBM peak as the
    main roofline.  Parity of both stages: tests/test_gpu_zt_transport.py (bitwise against the oracle up to the
    1-degree tripole; the 0.1-degree reconstruction holds 249 planes = 19 GB on the device, its oracle twin is not run
    here)."""
    from cice5_b200 import fixtures as fx
    peak, peak_src = measured_peaks()
    cells = grid.dom.nx * grid.dom.ny
    e.set_transport_grid(grid.f)
    dep = []
    for k in range(warmup + steps):
        e.departure_points(dt, True)
        if k >= warmup:
            dep.append(e.transport_timing())
    d = fx.transport_tracers(grid, case.inp, 1, "cice")
    ncat, nt = d["tm"].shape[0], d["tm"].shape[1]
    e.set_transport_tracers(ncat, d["tracer_type"], d["depend"], d["has_dependents"], d)
    e.construct_fields_upload(d["mm"], d["tm"])
    rec = []
    for k in range(warmup + steps):
        e.construct_fields_compute()
        if k >= warmup:
            rec.append(e.transport_timing())
    # all of horizontal_remap (reconstruction + departure points + their ghost updates + edge fluxes + update_fields);
    # mm / tm are updated in place, so every call starts from a fresh upload (not timed)
    rem = []
    stop = None
    for k in range(1 + 3):
        e.construct_fields_upload(d["mm"], d["tm"])
        e.horizontal_remap_compute(dt)
        if k >= 1:
            rem.append(e.transport_timing())
    mm_new, tm_new, stop = e.horizontal_remap_fetch()
    check = None
    if cpu:
        # the same horizontal_remap on one host core by the oracle port, from the velocities on the device: a parity leg
        # (every cell of mm, tm bitwise) and the CPU baseline of this stage
        from oracle import oracle as orc
        pio = {k: np.zeros_like(v) for k, v in case.io.items()}
        pout = {k: np.zeros_like(v) for k, v in case.out.items()}
        e.download(pio, pout)
        geo = dict(d)
        geo.update({k: grid.f[k] for k in ("dxu", "dyu", "HTN", "HTE", "tarear")})
        rmm, rtm = d["mm"].copy(), d["tm"].copy()
        t0 = time.perf_counter()
        rstop = orc.horizontal_remap(grid.dom, dt, d["tracer_type"], d["depend"], d["has_dependents"], geo, pio["uvel"],
                                     pio["vvel"], rmm, rtm)
        cpu_ms = 1e3 * (time.perf_counter() - t0)
        check = {"bitwise": bool(np.array_equal(mm_new, rmm) and np.array_equal(tm_new, rtm) and tuple(stop) == tuple(rstop)),
                 "arrays": "mm (ncat+1 planes), tm (ncat x ntrace planes), every cell; the stop report",
                 "cpu_port_ms": cpu_ms, "cpu_cores": 1, "cpu_kind": "port (oracle/remap_oracle.c, serial)"}
    med = lambda rows, key: float(np.median([r[key] for r in rows]))
    # algorithmic bytes per cell: departure reads u, v, dxu, dyu, HTN, HTE and writes dpx, dpy twice (first-order,
    # midpoint) -> 10 doubles; reconstruction, per ice category, reads mm + ntrace tracers + hm xav yav xxav yyav and
    # writes mc mx my + 3 ntrace planes; open water reads mm hm xav yav and writes mc mx my
    dep_bytes = 8 * 10 * cells
    rec_bytes = 8 * cells * (ncat * (6 + nt) + ncat * 3 * (1 + nt) + 4 + 3)
    dms, rms = med(dep, "departure_ms"), med(rec, "reconstruct_ms")
    return {"grid": grid.dom.name, "ncat": ncat, "ntrace": nt, "steps": steps, "warmup": warmup,
            "departure_points": {"kernel_ms": dms, "halo_ms": med(dep, "departure_halo_ms"), "alg_bytes": dep_bytes,
                                 "achieved_GBps": dep_bytes / (dms * 1e-3) / 1e9, "frac": dep_bytes / (dms * 1e-3) / 1e9 / peak},
            "construct_fields": {"kernel_ms": rms, "halo_ms": med(rec, "reconstruct_halo_ms"), "alg_bytes": rec_bytes,
                                 "achieved_GBps": rec_bytes / (rms * 1e-3) / 1e9, "frac": rec_bytes / (rms * 1e-3) / 1e9 / peak,
                                 "halo_updates": 3 * (ncat + 1) + 3 * ncat * nt},
            # per cell: the two edge sweeps gather dpx dpy dxu dyu + (mc mx my) x (ncat+1) + (tc tx ty) x ncat x ntrace and
            # write (ncat+1) + ncat x ntrace flux planes each; update_fields reads both flux sets, mm, tm, tarear and
            # writes mm, tm
            "horizontal_remap": (lambda fb, ms, fms: {"remap_ms": ms, "fluxes_and_update_ms": fms, "fluxes_and_update_alg_bytes": fb,
                                                      "fluxes_and_update_GBps": fb / (fms * 1e-3) / 1e9,
                                                      "fluxes_and_update_frac": fb / (fms * 1e-3) / 1e9 / peak,
                                                      "stop": list(stop), "max_abs_area_change": float(np.abs(mm_new - d["mm"])[:, 1:-1, 1:-1].max()),
                                                      "oracle_check": check})(
                8 * cells * (2 * (4 + 3 * (ncat + 1) + 3 * ncat * nt) + 2 * ((ncat + 1) + ncat * nt)        # edge sweeps
                             + 2 * ((ncat + 1) + ncat * nt) + 2 * ((ncat + 1) + ncat * nt) + 1),             # update_fields
                med(rem, "remap_ms"), med(rem, "flux_update_ms")),
            "peak_GBps": peak, "peak_source": peak_src,
            "parity": "tests/test_gpu_zt_transport.py: bitwise against oracle/remap_oracle.c (pinned to the reference text)"}


def eap_block(args, device=0):
    """EAP (kdyn = 2) beside the headline, N = 1: `eap(dt)` through the same entry points.
    (a) the 1-degree tripole, one whole call at ndte = 120: every inout / out field and every EAP field bit for bit against
        the oracle (outside any timed region), device time of the ndte loop;
    (b) the bench grid, ndte = 11 (stepa on subcycles 1 and 11), three calls: device time per subcycle (its parity at this
        size: tools/profile_eap.py, profiles/r02_eap.md).
    The lookup tables are an input of the path (the host's init_eap): fixtures.eap_tables()."""
    from cice5_b200 import evp, fixtures as fx
    tables = fx.eap_tables()

    def state(dom, seed=5):
        r = np.random.default_rng(seed)
        sh = (dom.ny + 2, dom.nx + 2)
        e = {}
        for c in (1, 2, 3, 4):                    # a valid structure tensor: eigenvalues inside (0, 1)
            e[f"a11_{c}"] = r.uniform(0.3, 0.9, sh)
            e[f"a12_{c}"] = r.uniform(-0.9, 0.9, sh) * np.sqrt(e[f"a11_{c}"] * (1.0 - e[f"a11_{c}"]))
        for n in ("a11", "a12") + evp.EAP_OUT:
            e[n] = np.zeros(sh)
        return e

    def run(gridname, ndte, calls, check):
        grid = fx.make_grid(gridname)
        dt = fx.GRID_DT.get(gridname, 3600.0)
        case = fx.make_case(grid, args.scenario, ndte=ndte, dt=dt)
        e0 = state(case.dom)
        got_e = {k: v.copy() for k, v in e0.items()}
        io = {k: v.copy() for k, v in case.io.items()}
        e = evp.EvpGpu.for_domain(grid.dom, ndte, dt, device=device, strict=True, kdyn=2)
        try:
            e.set_grid(grid.f)
            e.set_eap_tables(tables)
            e.upload_eap(got_e)
            ms = []
            out = None
            for k in range(calls):
                out = fx.zeroed_outputs(case.dom)
                e.evp(dt, case.inp, io, out)
                st = e.stats()
                ms.append(st.loop_ms)
                if k == 0 and check:
                    e.download_eap(got_e)
                    first = ({n: v.copy() for n, v in io.items()}, out, {n: v.copy() for n, v in got_e.items()})
            r = {"workload": workload_name(args, grid.dom, dt, gridname).replace(f"ndte={args.ndte}", f"ndte={ndte}") + " kdyn=2 (EAP)",
                 "ndte": ndte, "active_t_cells": int(st.active_t_cells), "loop_ms": min(ms),
                 "us_per_subcycle": 1e3 * min(ms) / ndte, "value": st.active_t_cells * ndte / (min(ms) * 1e-3),
                 "unit": "cell-subcycles/s", "loop_kernel_launches": int(st.loop_kernel_launches)}
            if check:
                from oracle import oracle as orc                     # the checker, outside every timed region
                ref = fx.make_case(grid, args.scenario, ndte=ndte, dt=dt)
                ref_e = {k: v.copy() for k, v in e0.items()}
                t0 = time.perf_counter()
                ost = orc.run_case_eap(ref, ref_e, tables, nthreads=len(os.sched_getaffinity(0)))
                cpu_s = time.perf_counter() - t0
                gio, gout, ge = first
                ph = lambda a: a[1:-1, 1:-1]
                bad = [n for n in PARITY_IO if not np.array_equal(ph(gio[n]), ph(ref.io[n]))]
                bad += [n for n in PARITY_OUT if gout.get(n) is not None and not np.array_equal(ph(gout[n]), ph(ref.out[n]))]
                bad += [n for n in evp.EAP_CORNER + evp.EAP_AVG + evp.EAP_OUT if not np.array_equal(ph(ge[n]), ph(ref_e[n]))]
                r["parity"] = {"bitwise": not bad, "fields": len(PARITY_IO) + len(PARITY_OUT) + 19, "differing": bad[:8],
                               "oracle": "oracle/eap_oracle.inc, pinned to the reference text (its atan2 / sin / cos: csrc/evp_trig.h)"}
                r["cpu_port"] = {"value": ost.active_t_cells * ndte / ost.loop_seconds, "unit": "cell-subcycles/s",
                                 "cores": len(os.sched_getaffinity(0)), "call_s": cpu_s}
            return r
        finally:
            e.close()

    blk = {"tp1": run("tp1", 120, 2, True)}
    if args.grid != "tp1":
        blk[args.grid] = run(args.grid, 11, 3, False)
    return blk


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args, int(os.environ.get("RANK", "0")))
        return 0
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("launch N > 1 with torch.distributed.run --nproc-per-node N")
    from cice5_b200 import evp, fixtures as fx
    if evp.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the EVP path has no CPU fallback")
    ctx = Ctx()
    rank, N = ctx.rank, ctx.world

    # CPU baseline first (rank 0, N = 1): its finished case is also the reference of the parity leg
    cpu, ref_case = None, None
    if rank == 0 and N == 1 and not args.no_cpu_baseline:
        try:
            grid = fx.make_grid(args.grid)
            dt = fx.GRID_DT.get(args.grid, 3600.0)
            kept = []
            cpu = cpu_sample(args, lambda n: fx.make_case(grid, args.scenario, ndte=n, dt=dt), keep=kept)
            if args.cpu_ndte == args.ndte:
                ref_case = kept[0]
        except Exception as ex:       # the oracle is only the reported baseline
            cpu = {"value": None, "error": str(ex)}

    traffic = NCU_TRAFFIC.get((args.grid, args.scenario, N, not args.no_compact_metrics))
    res = measure(ctx, args, args.grid, args.scenario, args.steps, args.warmup, not args.no_e2e,
                  not args.no_parity, ref_case, traffic)
    others = {}
    if not args.no_other_configs and args.grid == "tp01":
        # the other BASELINE grid shapes, short runs (not bench lines of their own: SURVEY 8d table)
        for g in ("gx3", "gx1", "tp1", "tp025"):
            if N > 1 and g in ("gx3", "gx1") and N > 2:
                continue                           # 100 / 320 columns: narrower than one strip per rank beyond 2 ranks
            try:
                r = measure(ctx, args, g, args.scenario, 10, 3, False, not args.no_parity, None, None)
            except Exception as ex:               # noqa: BLE001
                r = {"error": str(ex)[:200]} if rank == 0 else None
            if rank == 0:
                others[g] = r if "error" in r else {
                    "workload": r["config"]["workload"], "value": r["value"], "us_per_subcycle": r["us_per_subcycle"],
                    "ms_per_step": r["ms_per_step"], "roofline_frac": r["roofline"]["frac"],
                    "active_t_cells": r["config"]["active_t_cells"],
                    "regime": "HBM" if ALG_BYTES * r["roofline"]["active_cells_per_launch"] > L2_BYTES else "L2-resident",
                    "parity": r["parity"], "value_check": r["value_check"],
                    "loop_kernel_launches_per_step": r["loop_kernel_launches_per_step"]}
                if r.get("transport"):
                    others[g]["transport"] = r["transport"]
    if rank == 0 and N == 1 and not args.no_eap:
        try:
            res["eap"] = eap_block(args, ctx.local_rank)
        except Exception as ex:                   # noqa: BLE001 -- a side block must not cost the headline line
            res["eap"] = {"error": str(ex)[:300]}
    if rank == 0:
        if cpu is not None:
            res["cpu_baseline"] = cpu
        if others:
            res["other_configs"] = others
        print(json.dumps(res), flush=True)
    if ctx.dist is not None:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
