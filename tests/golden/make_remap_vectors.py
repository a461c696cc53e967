#!/usr/bin/env python
"""Golden vectors for the first stage of horizontal transport from the REFERENCE'S OWN TEXT
(run in the build container only): `departure_points` of source/ice_transport_remap.F90:1493-1670, executed
mechanically by tests/golden/fortran_exec.py on seeded blocks -- first-order and midpoint-corrected departure points,
cells at rest, and a block whose velocities send a departure point more than one cell away (l_stop, istop, jstop).

    python tests/golden/make_remap_vectors.py        # rewrites tests/golden/ref_remap_vectors.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import fortran_exec as fe  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(HERE, "ref_remap_vectors.npz")
NXB, NYB = 14, 11            # block incl. the ghost ring


def load():
    def lines(rel):
        return fe.preprocess(open(os.path.join(REF, rel)).read(), ())
    ns = {}
    fe.parse_parameters(lines("drivers/cice/ice_constants.F90"), ns)
    ns.update(nu_diag=6, my_task=0)
    m = fe.Module(ns)
    m.load(lines("source/ice_transport_remap.F90"), ["departure_points"])
    return m


def block(seed, speed):
    r = np.random.default_rng(seed)
    sh = (NYB, NXB)
    d = {"uvel": r.uniform(-speed, speed, sh), "vvel": r.uniform(-speed, speed, sh),
         "dxu": r.uniform(2.0e4, 3.0e4, sh), "dyu": r.uniform(2.0e4, 3.0e4, sh),
         "HTN": r.uniform(2.0e4, 3.0e4, sh), "HTE": r.uniform(2.0e4, 3.0e4, sh)}
    rest = r.uniform(0, 1, sh) < 0.2                     # cells at rest: the midpoint branch is skipped
    d["uvel"][rest] = 0.0
    d["vvel"][rest] = 0.0
    half = r.uniform(0, 1, sh) < 0.1                     # one component zero
    d["vvel"][half] = 0.0
    return d


def run(m, d, dt, midpt):
    """One call; arrays are (NYB, NXB) C order == Fortran (nx_block, ny_block).  The out arrays start with garbage:
    the routine zeroes them itself (:1557-1558)."""
    a = {k.lower(): fe.FArr.from_numpy(v) for k, v in d.items()}
    dpx, dpy = fe.FArr.from_numpy(np.full((NYB, NXB), -7.0)), fe.FArr.from_numpy(np.full((NYB, NXB), -7.0))
    r = m.call("departure_points", nx_block=NXB, ny_block=NYB, ilo=2, ihi=NXB - 1, jlo=2, jhi=NYB - 1, nghost=1, dt=dt,
               dpx=dpx, dpy=dpy, l_dp_midpt=midpt, l_stop=False, istop=0, jstop=0, **a)
    return {"dpx": dpx.to_numpy(), "dpy": dpy.to_numpy(), "l_stop": int(bool(r["l_stop"])), "istop": int(r["istop"]),
            "jstop": int(r["jstop"])}


def main():
    m = load()
    out = {"nxb": NXB, "nyb": NYB}
    cases = [("slow_mid", 11, 0.6, 1), ("slow_first", 12, 0.6, 0), ("fast_mid", 13, 2.5, 1), ("stop", 14, 12.0, 1)]
    for name, seed, speed, midpt in cases:
        d = block(seed, speed)
        dt = 3600.0
        res = run(m, d, dt, bool(midpt))
        for k, v in d.items():
            out[f"{name}/{k}"] = v
        out[f"{name}/dt"] = dt
        out[f"{name}/l_dp_midpt"] = midpt
        for k, v in res.items():
            out[f"{name}/out/{k}"] = v
        print(name, "l_stop", res["l_stop"], "istop, jstop", res["istop"], res["jstop"],
              "max |dpx|", float(np.abs(res["dpx"]).max()))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
