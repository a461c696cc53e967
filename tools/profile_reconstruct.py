#!/usr/bin/env python
"""One field reconstruction (evp_gpu_construct_fields_compute) on a bench grid, for ncu / timing runs:
    python tools/profile_reconstruct.py [grid] [calls]
Prints the device time of each call (evp_gpu_get_transport_timing)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cice5_b200 import evp, fixtures as fx  # noqa: E402


def main():
    gridname = sys.argv[1] if len(sys.argv) > 1 else "tp01"
    calls = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    grid = fx.make_grid(gridname)
    dt = fx.GRID_DT.get(gridname, 3600.0)
    case = fx.make_case(grid, "full", ndte=120, dt=dt)
    d = fx.transport_tracers(grid, case.inp, 1, "cice")
    e = evp.EvpGpu.for_domain(grid.dom, 120, dt, strict=True)
    try:
        e.set_transport_tracers(d["tm"].shape[0], d["tracer_type"], d["depend"], d["has_dependents"], d)
        e.construct_fields_upload(d["mm"], d["tm"])
        for k in range(calls):
            e.construct_fields_compute()
            t = e.transport_timing()
            print(f"call {k}: reconstruct {t['reconstruct_ms']:.3f} ms, ghost updates {t['reconstruct_halo_ms']:.3f} ms", flush=True)
    finally:
        e.close()


if __name__ == "__main__":
    main()
