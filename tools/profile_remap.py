#!/usr/bin/env python
"""One evp call, then horizontal_remap (evp_gpu_horizontal_remap_compute) on a bench grid, for timing / ncu runs:
    python tools/profile_remap.py [grid] [calls] [dt_scale] [scenario]
Prints the device time of each stage (evp_gpu_get_transport_timing)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cice5_b200 import evp, fixtures as fx  # noqa: E402


def main():
    gridname = sys.argv[1] if len(sys.argv) > 1 else "tp01"
    calls = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    dt_scale = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    scenario = sys.argv[4] if len(sys.argv) > 4 else "full"
    grid = fx.make_grid(gridname)
    dt = fx.GRID_DT.get(gridname, 3600.0)
    case = fx.make_case(grid, scenario, ndte=120, dt=dt)
    d = fx.transport_tracers(grid, case.inp, 1, "cice")
    e = evp.EvpGpu.for_domain(grid.dom, 120, dt, strict=True)
    try:
        e.set_grid(grid.f)
        e.set_transport_grid(grid.f)
        e.set_transport_tracers(d["tm"].shape[0], d["tracer_type"], d["depend"], d["has_dependents"], d)
        e.evp(dt, case.inp, case.io, case.out)
        print(f"evp done, umax {max(np.abs(case.io['uvel']).max(), np.abs(case.io['vvel']).max()):.3f} m/s", flush=True)
        for k in range(calls):
            e.construct_fields_upload(d["mm"], d["tm"])
            e.horizontal_remap_compute(dt * dt_scale)
            t = e.transport_timing()
            print(f"call {k}: remap {t['remap_ms']:.3f} ms = reconstruct {t['reconstruct_ms']:.3f} + ghost updates "
                  f"{t['reconstruct_halo_ms']:.3f} + departure {t['departure_ms']:.3f} + ghost updates {t['departure_halo_ms']:.3f}"
                  f" + fluxes and update {t['flux_update_ms']:.3f}", flush=True)
        mm, tm, stop = e.horizontal_remap_fetch()
        print("stop", stop, "max |d mm|", float(np.abs(mm - d["mm"])[:, 1:-1, 1:-1].max()), flush=True)
    finally:
        e.close()


if __name__ == "__main__":
    main()
