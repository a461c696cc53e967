"""EAP (kdyn = 2) with one PROCESS per GPU: every rank runs eap(dt) on its x-slab with peer-memory halos (CUDA IPC, the
cross-rank flag barrier between the stages of a subcycle), rank 0 gathers the slabs and compares them BITWISE with the
single-block oracle.  No torch: the ranks are `multiprocessing` children, the peer blobs travel through queues.

    python tests/run_multigpu_eap.py --nranks 2 --grid tp1 --scenario caps --ndte 21 --calls 2
"""
import argparse
import multiprocessing as mp
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(rank, world, args, blob_q, blob_back, result_q):
    import numpy as np
    from cice5_b200 import evp, fixtures as fx
    from tests.eap_util import eap_state
    from oracle import oracle as orc                       # rank-local: only for init_eap's tables (host code)
    tables = orc.eap_tables()
    grid = fx.make_grid(args.grid)
    case = fx.make_case(grid, args.scenario, ndte=args.ndte)
    dom = case.dom
    starts = fx.balanced_slab_starts(grid.f["tmask"][1:-1, 1:-1] > 0, world) if args.balanced else None
    sl = lambda d: {k: fx.slab(v, rank, world, starts) for k, v in d.items()}
    e = evp.EvpGpu.for_domain(dom, args.ndte, case.dt, rank=rank, nranks=world, device=rank, strict=True,
                              use_cuda_graph=bool(args.graph), slab_start=starts, kdyn=2)
    e.set_grid(sl({k: grid.f[k] for k in evp.GRID_R8 + ("tmask", "umask", "HTE", "HTN")}))
    blob_q.put((rank, e.peer_blob()))
    e.set_peers(blob_back.get())
    e.set_eap_tables(tables)
    eap = sl(eap_state(dom))
    e.upload_eap(eap)
    inp, io = sl(case.inp), sl(case.io)
    for _ in range(args.calls):
        out = sl(fx.zeroed_outputs(dom))
        e.evp(case.dt, inp, io, out)
    e.download_eap(eap)
    st = e.stats()
    result_q.put((rank, io, out, eap, st.loop_ms, st.kernel_launches))
    e.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nranks", type=int, default=2)
    ap.add_argument("--grid", default="tp1")
    ap.add_argument("--scenario", default="caps")
    ap.add_argument("--ndte", type=int, default=21)
    ap.add_argument("--calls", type=int, default=2)
    ap.add_argument("--graph", type=int, default=1)
    ap.add_argument("--balanced", type=int, default=1)
    args = ap.parse_args()
    world = args.nranks
    ctx = mp.get_context("spawn")
    blob_q, result_q = ctx.Queue(), ctx.Queue()
    backs = [ctx.Queue() for _ in range(world)]
    procs = [ctx.Process(target=worker, args=(r, world, args, blob_q, backs[r], result_q)) for r in range(world)]
    for p in procs:
        p.start()
    blobs = dict(blob_q.get(timeout=300) for _ in range(world))
    for q in backs:
        q.put([blobs[r] for r in range(world)])
    res = dict((r[0], r[1:]) for r in (result_q.get(timeout=600) for _ in range(world)))
    for p in procs:
        p.join(timeout=60)

    import numpy as np
    from cice5_b200 import fixtures as fx
    from oracle import oracle as orc
    from tests.eap_util import clone_eap, compare_eap, eap_state
    from tests.util import clone_case
    grid = fx.make_grid(args.grid)
    case = fx.make_case(grid, args.scenario, ndte=args.ndte)
    tables = orc.eap_tables()
    ref, ref_e = clone_case(case), clone_eap(eap_state(case.dom))
    for _ in range(args.calls):
        ref.out = fx.zeroed_outputs(case.dom)
        orc.run_case_eap(ref, ref_e, tables)
    got = clone_case(case)
    got.io = {k: fx.unslab([res[r][0][k] for r in range(world)]) for k in res[0][0]}
    got.out = {k: fx.unslab([res[r][1][k] for r in range(world)]) for k in res[0][1]}
    got_e = {k: fx.unslab([res[r][2][k] for r in range(world)]) for k in res[0][2]}
    try:
        compare_eap(got, ref, got_e, ref_e, case.dom)
        print(f"MULTIGPU EAP PARITY OK: {args.grid} {args.scenario} N={world} processes, ndte={args.ndte}, calls={args.calls}: "
              f"bitwise == oracle on every field; loop {max(res[r][3] for r in range(world)):.3f} ms (slowest rank), "
              f"launches {res[0][4]}", flush=True)
        return 0
    except AssertionError as ex:
        print("MULTIGPU EAP PARITY FAILED:", str(ex)[:2000], flush=True)
        return 1


if __name__ == "__main__":
    sys.exit(main())
