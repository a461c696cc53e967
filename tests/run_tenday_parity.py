#!/usr/bin/env python
"""Ten-day parity run (SURVEY.md section 8d "Parity runs" (2); BASELINE.json north_star: "max and
RMS error reported over a 10-day run").

240 consecutive evp(dt) calls (dt = 3600 s), ice state fixed, forcing evolving (the cyclones cross
the domain once per 10 days), stresses / velocity / iceumask carried from call to call -- on the
device through the C ABI and on the CPU oracle, same bytes in.  After every call the inout fields
are compared on physical cells:
    err_inf = max|x - ref| / max|ref|      err_rms = rms(x - ref) / rms(ref)
and the report holds, per field, the maximum over the calls and the value after the last call, for
the product build (-fmad=false: expected 0, bitwise) and for the FMA-contracted comparison build.

    python tests/run_tenday_parity.py --grid gx3 --calls 240 --out gpurun_out/tenday_gx3.json
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

from cice5_b200 import evp, fixtures as fx
from tests.util import STATE, clone_case, phys, rel_errors

OUT_REPORT = ("strocnxT", "strocnyT", "divu", "shear", "strintx", "strinty")


def ten_day_run(grid_name: str, scenario: str, calls: int, ndte: int = 120, builds=("strict", "fma"),
                progress=None):
    """Returns {build: {field: {"max_inf", "max_rms", "final_inf", "final_rms"}}, ...}."""
    from oracle import oracle as orc
    grid = fx.make_grid(grid_name)
    case = fx.make_case(grid, scenario, ndte=ndte)
    ref = clone_case(case)
    runs = {b: clone_case(case) for b in builds}
    solvers = {}
    rep = {b: {} for b in builds}
    bitwise = {b: True for b in builds}
    try:
        for b in builds:
            e = evp.EvpGpu.for_domain(case.dom, ndte, case.dt, strict=(b == "strict"))
            e.set_grid(grid.f)
            solvers[b] = e
        for step in range(calls):
            f = fx.forcing(grid, case.inp, step, steps_per_cycle=240)
            ref.inp.update({k: v.copy() for k, v in f.items()})
            ref.out = fx.zeroed_outputs(case.dom)
            orc.run_case(ref)
            for b, c in runs.items():
                c.inp.update({k: v.copy() for k, v in f.items()})
                c.out = fx.zeroed_outputs(case.dom)
                solvers[b].evp(case.dt, c.inp, c.io, c.out)
                for group, names in (("io", STATE), ("out", OUT_REPORT)):
                    for n in names:
                        x, r = getattr(c, group)[n], getattr(ref, group)[n]
                        if not np.array_equal(phys(x), phys(r)):
                            bitwise[b] = False
                        ei, er = rel_errors(x, r)
                        d = rep[b].setdefault(n, {"max_inf": 0.0, "max_rms": 0.0})
                        d["max_inf"] = max(d["max_inf"], float(ei))
                        d["max_rms"] = max(d["max_rms"], float(er))
                        d["final_inf"], d["final_rms"] = float(ei), float(er)
                if not np.array_equal(phys(c.io["iceumask"]), phys(ref.io["iceumask"])):
                    bitwise[b] = False
            if progress:
                progress(step)
    finally:
        for e in solvers.values():
            e.close()
    umax = float(np.abs(phys(ref.io["uvel"])).max())
    return {"grid": grid_name, "scenario": scenario, "calls": calls, "ndte": ndte, "dt": case.dt,
            "oracle_umax_final": umax, "bitwise_all_calls": bitwise, "fields": rep}


def summarise(rep):
    lines = []
    for b, fields in rep["fields"].items():
        vel = max(fields[n]["max_inf"] for n in ("uvel", "vvel"))
        st = max(fields[n]["max_inf"] for n in fx.STRESS_FIELDS)
        vel_r = max(fields[n]["max_rms"] for n in ("uvel", "vvel"))
        st_r = max(fields[n]["max_rms"] for n in fx.STRESS_FIELDS)
        lines.append(f"{rep['grid']:5s} {rep['scenario']:5s} {rep['calls']:4d} calls  build={b:6s} "
                     f"bitwise={rep['bitwise_all_calls'][b]!s:5s} max-over-calls: vel inf {vel:.2e} rms {vel_r:.2e}"
                     f" | stress inf {st:.2e} rms {st_r:.2e}")
    return "\n".join(lines)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", default="gx3")
    ap.add_argument("--scenario", default="caps")
    ap.add_argument("--calls", type=int, default=240)
    ap.add_argument("--ndte", type=int, default=120)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    if evp.device_count() < 1:
        raise SystemExit("no CUDA device: the EVP path has no CPU fallback")
    t0 = time.perf_counter()
    rep = ten_day_run(args.grid, args.scenario, args.calls, args.ndte)
    rep["wall_s"] = time.perf_counter() - t0
    print(summarise(rep))
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(rep, f, indent=1)
    return 0 if rep["bitwise_all_calls"].get("strict", True) else 1


if __name__ == "__main__":
    sys.exit(main())
