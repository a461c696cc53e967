#!/usr/bin/env python
"""EAP (kdyn = 2) on a bench grid, for timing / parity / ncu runs:
    python tools/profile_eap.py [grid] [ndte] [calls] [check]
check = 1: the first call is compared with the CPU oracle bit for bit (every inout / out field and every EAP field).
Prints the device time of the ndte loop per call (evp_gpu_get_stats) and the oracle's time for the same call."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cice5_b200 import evp, fixtures as fx  # noqa: E402


def main():
    gridname = sys.argv[1] if len(sys.argv) > 1 else "tp01"
    ndte = int(sys.argv[2]) if len(sys.argv) > 2 else 11
    calls = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    check = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    from oracle import oracle as orc               # the checker (check = 1); the tables are fixtures' (an input of the path)
    from tests.eap_util import clone_eap, compare_eap, eap_state
    from tests.util import clone_case
    tables = fx.eap_tables()
    grid = fx.make_grid(gridname)
    dt = fx.GRID_DT.get(gridname, 3600.0)
    case = fx.make_case(grid, "full", ndte=ndte, dt=dt)
    e0 = eap_state(case.dom)
    got, got_e = clone_case(case), clone_eap(e0)
    e = evp.EvpGpu.for_domain(grid.dom, ndte, dt, strict=True, kdyn=2)
    try:
        e.set_grid(grid.f)
        e.set_eap_tables(tables)
        e.upload_eap(got_e)
        for k in range(calls):
            got.out = fx.zeroed_outputs(case.dom)
            e.evp(dt, got.inp, got.io, got.out)
            st = e.stats()
            print(f"call {k}: {gridname} ndte={ndte} EAP loop {st.loop_ms:.3f} ms = {1e3 * st.loop_ms / ndte:.1f} us per subcycle "
                  f"(incl. {(ndte + 9) // 10} stepa), {st.active_t_cells} T cells, "
                  f"{st.active_t_cells * ndte / (st.loop_ms * 1e-3):.3e} cell-subcycles/s, launches in the loop "
                  f"{st.loop_kernel_launches}", flush=True)
            if k == 0 and check:
                e.download_eap(got_e)
                ref, ref_e = clone_case(case), clone_eap(e0)
                t0 = time.time()
                stats = orc.run_case_eap(ref, ref_e, tables)
                print(f"oracle: {time.time() - t0:.1f} s for the call, loop {stats.loop_seconds:.2f} s = "
                      f"{stats.active_t_cells * ndte / stats.loop_seconds:.3e} cell-subcycles/s "
                      f"({os.cpu_count()} host threads)", flush=True)
                compare_eap(got, ref, got_e, ref_e, case.dom)
                print("device == oracle bit for bit on every field (state, outputs, structure tensor, diagnostics)", flush=True)
    finally:
        e.close()


if __name__ == "__main__":
    main()
