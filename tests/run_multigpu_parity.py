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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", default="tp1")
    ap.add_argument("--scenario", default="caps")
    ap.add_argument("--ndte", type=int, default=120)
    ap.add_argument("--calls", type=int, default=2)
    ap.add_argument("--graph", type=int, default=1)
    ap.add_argument("--balanced", type=int, default=1, help="slabs balanced by ocean cells")
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
    for call in range(args.calls):
        out = sl(fx.zeroed_outputs(dom))
        e.evp(case.dt, inp, io, out)
    st = e.stats()
    gathered = [None] * world
    dist.gather_object((io, out), gathered if rank == 0 else None, dst=0)
    ok = True
    if rank == 0:
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
    e.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
