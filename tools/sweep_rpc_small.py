#!/usr/bin/env python
"""Rows per work item (EVP_B200_RPC) of the persistent loop kernel on the L2-resident grids: device time of the ndte loop
per subcycle for each setting (classic EVP, ndte = 120, one GPU).  python tools/sweep_rpc_small.py [grids] [rpcs]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cice5_b200 import evp, fixtures as fx  # noqa: E402


def main():
    grids = sys.argv[1].split(",") if len(sys.argv) > 1 else ["gx3", "gx1", "tp1"]
    rpcs = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64]
    for g in grids:
        grid = fx.make_grid(g)
        dt = fx.GRID_DT.get(g, 3600.0)
        case = fx.make_case(grid, "full", ndte=120, dt=dt)
        row = []
        for rpc in rpcs:
            if rpc:
                os.environ["EVP_B200_RPC"] = str(rpc)
            else:
                os.environ.pop("EVP_B200_RPC", None)          # 0 = the library's own rule
            e = evp.EvpGpu.for_domain(grid.dom, 120, dt, strict=True, use_cuda_graph=True)
            try:
                e.set_grid(grid.f)
                e.upload(case.inp, case.io)
                best = 1e9
                for _ in range(4):
                    e.compute(dt)
                    best = min(best, e.stats().loop_ms)
                row.append((rpc, 1e3 * best / 120))
            finally:
                e.close()
        print(g, "us per subcycle by rows per item: " + ", ".join(f"{r or 'rule'}: {t:.2f}" for r, t in row), flush=True)


if __name__ == "__main__":
    main()
