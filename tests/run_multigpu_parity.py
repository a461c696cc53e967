"""One process per GPU (torchrun): every rank runs evp(dt) on its x-slab with peer-memory halos
(CUDA IPC), rank 0 gathers the slabs and compares them BITWISE with the single-block oracle.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tests/run_multigpu_parity.py --grid tp1 --calls 2
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

from cice5_b200 import evp, fixtures as fx


def evolve_run(args, e, grid, case, sl, rank, world, starts):
    """The 10-day parity run on N processes: forcing of call `step` on every rank, state carried on the device
    slabs and in the oracle, every inout field + strocnxT/yT, divu, shear compared after EVERY call."""
    import json
    import time
    from tests.util import STATE, clone_case, phys, rel_errors
    dom = case.dom
    report = ("strocnxT", "strocnyT", "divu", "shear")
    io = sl(case.io)
    ref = clone_case(case) if rank == 0 else None
    if rank == 0:
        from oracle import oracle as orc
    stats = {}
    bitwise = True
    t0 = time.perf_counter()
    for step in range(args.calls):
        f = fx.forcing(grid, case.inp, step, steps_per_cycle=240)
        inp = dict(case.inp)
        inp.update(f)
        out = sl(fx.zeroed_outputs(dom))
        e.evp(case.dt, sl(inp), io, out)
        gathered = [None] * world
        small = ({k: io[k] for k in STATE + ("iceumask",)}, {k: out[k] for k in report})
        dist.gather_object(small, gathered if rank == 0 else None, dst=0)
        if rank == 0:
            ref.inp.update({k: v.copy() for k, v in f.items()})
            ref.out = fx.zeroed_outputs(dom)
            orc.run_case(ref)
            for group, names in ((0, STATE), (1, report)):
                for n in names:
                    x = fx.unslab([g[group][n] for g in gathered])
                    r = (ref.io if group == 0 else ref.out)[n]
                    if not np.array_equal(phys(x), phys(r)):
                        bitwise = False
                    ei, er = rel_errors(x, r)
                    d = stats.setdefault(n, {"max_inf": 0.0, "max_rms": 0.0})
                    d["max_inf"] = max(d["max_inf"], float(ei)); d["max_rms"] = max(d["max_rms"], float(er))
                    d["final_inf"], d["final_rms"] = float(ei), float(er)
            if not np.array_equal(phys(fx.unslab([g[0]["iceumask"] for g in gathered])), phys(ref.io["iceumask"])):
                bitwise = False
    ok = True
    if rank == 0:
        vel = max(stats[n]["max_inf"] for n in ("uvel", "vvel"))
        velr = max(stats[n]["max_rms"] for n in ("uvel", "vvel"))
        st = max(stats[n]["max_inf"] for n in fx.STRESS_FIELDS)
        str_ = max(stats[n]["max_rms"] for n in fx.STRESS_FIELDS)
        rep = {"grid": args.grid, "scenario": args.scenario, "calls": args.calls, "ndte": args.ndte, "nranks": world,
               "processes": world, "slab_starts": starts, "bitwise_all_calls": bitwise,
               "vel_max_inf": vel, "vel_max_rms": velr, "stress_max_inf": st, "stress_max_rms": str_,
               "vel_final_inf": max(stats[n]["final_inf"] for n in ("uvel", "vvel")),
               "oracle_umax_final": float(np.abs(phys(ref.io["uvel"])).max()), "wall_s": time.perf_counter() - t0,
               "fields": stats}
        print("TENDAY " + json.dumps({k: v for k, v in rep.items() if k != "fields"}), flush=True)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as fh:
                json.dump(rep, fh, indent=1)
        ok = bitwise
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", default="tp1")
    ap.add_argument("--scenario", default="caps")
    ap.add_argument("--ndte", type=int, default=120)
    ap.add_argument("--calls", type=int, default=2)
    ap.add_argument("--graph", type=int, default=1)
    ap.add_argument("--balanced", type=int, default=1, help="slabs balanced by ocean cells")
    ap.add_argument("--evolve", type=int, default=0,
                    help="1: the forcing evolves from call to call (the 10-day run of SURVEY 8d: 240 calls) and EVERY "
                         "call is compared; prints max / final err_inf and err_rms per field group")
    ap.add_argument("--transport", type=int, default=0,
                    help="1: after the evp calls, transport_remap on every rank's slab from the resident velocities "
                         "(peer-memory halos of the transport stages), compared bitwise with the oracle's; 2: three "
                         "calls, device times only (grids whose oracle twin does not fit the host)")
    ap.add_argument("--no-evp-check", type=int, default=0, help="1: skip the oracle comparison of the evp calls")
    ap.add_argument("--dt-scale", type=float, default=6.0, help="transport time step = dt * this")
    ap.add_argument("--out", default=None, help="JSON report (rank 0)")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    grid = fx.make_grid(args.grid)
    case = fx.make_case(grid, args.scenario, ndte=args.ndte)
    dom = case.dom
    starts = None
    if args.balanced:
        starts = fx.balanced_slab_starts(grid.f["tmask"][1:-1, 1:-1] > 0, world)
    sl = lambda d: {k: fx.slab(v, rank, world, starts) for k, v in d.items()}
    e = evp.EvpGpu.for_domain(dom, args.ndte, case.dt, rank=rank, nranks=world, device=local,
                              strict=True, use_cuda_graph=bool(args.graph), slab_start=starts)
    e.set_grid(sl({k: grid.f[k] for k in evp.GRID_R8 + ("tmask", "umask", "HTE", "HTN")}))
    blobs = [None] * world
    dist.all_gather_object(blobs, e.peer_blob())
    e.set_peers(blobs)
    inp, io = sl(case.inp), sl(case.io)
    if args.evolve:
        ok = evolve_run(args, e, grid, case, sl, rank, world, starts)
        e.close()
        dist.barrier()
        dist.destroy_process_group()
        sys.exit(0 if ok else 1)
    for call in range(args.calls):
        out = sl(fx.zeroed_outputs(dom))
        e.evp(case.dt, inp, io, out)
    st = e.stats()
    gathered = [None] * world
    dist.gather_object((io, out), gathered if rank == 0 else None, dst=0)
    ok = True
    tgath = None
    if args.transport:
        from tests.test_emu_device_sources import RHOS, LFRESH, _transport_state
        d, state = _transport_state(grid, case, 9)
        slab = lambda a: fx.slab(np.ascontiguousarray(a), rank, world, starts)
        e.set_transport_grid(sl({k: grid.f[k] for k in ("dxu", "dyu", "HTN", "HTE")}))
        e.set_transport_tracers(state["aicen"].shape[0], d["tracer_type"], d["depend"], d["has_dependents"],
                                sl({k: d[k] for k in ("hm", "xav", "yav", "xxav", "yyav")}))
        dist.barrier()
        mine = {k: slab(v) for k, v in state.items()}
        for rep in range(3 if args.transport == 2 else 1):
            e.transport_remap_upload(mine)
            dist.barrier()                        # every rank has its state on the device: the timed part starts together
            e.transport_remap_compute(case.dt * args.dt_scale, 6, 1, RHOS, LFRESH)
            new, stop = e.transport_remap_fetch()
        tms = e.transport_timing()
        if args.transport == 2:
            new = None
        tgath = [None] * world
        dist.gather_object((new, stop, tms), tgath if rank == 0 else None, dst=0)
    if rank == 0 and args.no_evp_check:
        print(f"MULTIGPU EVP (not compared): {args.grid} N={world} loop {st.loop_ms:.3f} ms", flush=True)
    if rank == 0 and not args.no_evp_check:
        from oracle import oracle as orc
        from tests.util import clone_case, compare_cases
        from tests.test_gpu_parity import OUT_CHECK
        ref = clone_case(case)
        for call in range(args.calls):
            ref.out = fx.zeroed_outputs(dom)
            orc.run_case(ref)
        got = clone_case(case)
        got.io = {k: fx.unslab([g[0][k] for g in gathered]) for k in io}
        got.out = {k: fx.unslab([g[1][k] for g in gathered]) for k in out}
        try:
            compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
            print(f"MULTIGPU PARITY OK: {args.grid} {args.scenario} N={world} calls={args.calls} "
                  f"bitwise == oracle; loop {st.loop_ms:.3f} ms, launches {st.kernel_launches}", flush=True)
        except AssertionError as ex:
            ok = False
            print("MULTIGPU PARITY FAILED:", str(ex)[:2000], flush=True)
    if rank == 0 and args.transport == 2:
        slow = {k: max(g[2][k] for g in tgath) for k in tgath[0][2]}
        print(f"MULTIGPU TRANSPORT TIMING: {args.grid} N={world} stops {[tuple(g[1]) for g in tgath]} slowest rank, ms: "
              + ", ".join(f"{k} {v:.3f}" for k, v in slow.items()), flush=True)
    if rank == 0 and args.transport == 1:
        geo = dict(d)
        geo.update({k: grid.f[k] for k in ("dxu", "dyu", "HTN", "HTE", "tarear")})
        rs = {k: v.copy() for k, v in state.items()}
        rstop = orc.transport_remap(dom, case.dt * args.dt_scale, d["tracer_type"], d["depend"], d["has_dependents"], geo,
                                    ref.io["uvel"], ref.io["vvel"], rs, 6, 1, RHOS, LFRESH)
        bad = [k for k in rs if not np.array_equal(fx.unslab([g[0][k] for g in tgath]), rs[k])]
        stops = [tuple(g[1]) for g in tgath]
        tok = not bad and all(s == tuple(rstop) for s in stops)
        moved = float(np.abs(rs["aicen"] - state["aicen"])[:, 1:-1, 1:-1].max())
        print(f"MULTIGPU TRANSPORT {'OK' if tok else 'FAILED'}: {args.grid} N={world} transport_remap bitwise == oracle "
              f"(ghost cells included), stops {stops} oracle {rstop}, mismatching arrays {bad}, max |d aicen| {moved:.3e}, "
              f"remap {max(g[2]['remap_ms'] for g in tgath):.3f} ms (slowest rank)", flush=True)
        ok = ok and tok
    e.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
