#!/usr/bin/env python
"""Random small domains, tracer tables and ice states through the host build of the transport device sources (tests/emu)
against the CPU oracle, bit for bit: evp, then horizontal_remap from the resident velocities -- tripole / displaced-pole /
closed-basin grids of 4..78 x 4..29 cells, 1-4 ranks, 1-3 ice categories, 0-9 tracers with random (valid) dependency
tables of all three types, init_remap or perturbed geometry, transport steps from a hundredth of a cell to most of one
(and beyond: the stop reports must agree too).

    python tools/fuzz_transport_on_emulator.py SEED [CASES]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from cice5_b200 import fixtures as fx
from oracle import oracle as orc
from tests.emu import emu
from tests.util import clone_case

rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
cases = int(sys.argv[2]) if len(sys.argv) > 2 else 30


def tracer_table(nt):
    """A random valid table: type 2 on an earlier type-1 tracer, type 3 on an earlier type-2 tracer."""
    tt, dep = [], []
    for k in range(nt):
        ones = [i for i, t in enumerate(tt) if t == 1]
        twos = [i for i, t in enumerate(tt) if t == 2]
        choice = rng.random()
        if choice < 0.3 and twos:
            tt.append(3); dep.append(int(rng.choice(twos)) + 1)
        elif choice < 0.7 and ones:
            tt.append(2); dep.append(int(rng.choice(ones)) + 1)
        else:
            tt.append(1); dep.append(0)
    hd = [0] * nt
    for d in dep:
        if d:
            hd[d - 1] = 1
    return np.array(tt, dtype=np.int32), np.array(dep, dtype=np.int32), np.array(hd, dtype=np.int32)


n_ok = n_bad = n_stop = 0
with emu.patched(kernel="tma") as evp:
    for it in range(cases):
        tripole = rng.random() < 0.5
        nx = int(rng.integers(2, 40)) * 2 if tripole else int(rng.integers(4, 70))
        ny = int(rng.integers(4, 30))
        info = dict(tripole=tripole, nx=nx, ny=ny)
        try:
            closed = bool(not tripole and rng.random() < 0.4)
            grid = fx.synthetic_tripole(nx, ny, wet_poles=bool(rng.random() < 0.5)) if tripole else (
                fx.synthetic_closed(nx, ny) if closed else fx.synthetic_displaced(nx, ny))
            case = fx.make_case(grid, ["full", "caps"][int(rng.integers(0, 2))], ndte=int(rng.integers(1, 5)), dt=3600.0)
            ncat, nt = int(rng.integers(1, 4)), int(rng.integers(0, 10))
            tt, dep, hd = tracer_table(nt)
            sh = case.inp["aice"].shape
            hm = (grid.f["tmask"] != 0).astype(np.float64)
            aicen = rng.uniform(0.0, 0.9 / ncat, (ncat,) + sh) * (rng.uniform(0, 1, (ncat,) + sh) < 0.7) * hm
            aicen[rng.uniform(0, 1, aicen.shape) < 0.05] = 3.0e-12
            mm = np.concatenate([(1.0 - aicen.sum(axis=0))[None] * hm, aicen])
            tm = rng.uniform(-2.0, 2.0, (ncat, nt) + sh) * (rng.uniform(0, 1, (ncat, nt) + sh) < 0.8)
            tm[rng.uniform(0, 1, tm.shape) < 0.03] = 2.0e-12
            geo = {"hm": hm}
            if rng.random() < 0.5:
                geo.update(xav=np.zeros(sh), yav=np.zeros(sh), xxav=np.full(sh, 1 / 12), yyav=np.full(sh, 1 / 12))
            else:
                geo.update(xav=rng.uniform(-0.04, 0.04, sh), yav=rng.uniform(-0.04, 0.04, sh),
                           xxav=rng.uniform(0.07, 0.1, sh), yyav=rng.uniform(0.07, 0.1, sh))
            dt = case.dt * float(10 ** rng.uniform(0.0, 2.6))
            ranks = [r for r in (1, 2, 3, 4) if nx % r == 0 and nx // r >= 2]
            nr = int(rng.choice(ranks))
            info.update(ncat=ncat, nt=nt, types=tt.tolist(), depend=dep.tolist(), nranks=nr, dt=round(dt), closed=closed)
            got = clone_case(case)
            f = grid.f
            if nr == 1:
                e = evp.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
                try:
                    e.set_grid(f); e.set_transport_grid(f); e.set_transport_tracers(ncat, tt, dep, hd, geo)
                    e.evp(case.dt, got.inp, got.io, got.out)
                    nmm, ntm, stop = e.horizontal_remap(dt, mm, tm)
                    stops = [stop]
                finally:
                    e.close()
            else:
                g = evp.EvpGroup(case.dom, case.ndte, case.dt, nr, strict=True)
                try:
                    g.set_grid(f); g.set_transport_grid(f); g.set_transport_tracers(ncat, tt, dep, hd, geo)
                    g.evp(case.dt, got.inp, got.io, got.out)
                    nmm, ntm, stops = g.horizontal_remap(dt, mm, tm)
                finally:
                    g.close()
            full = dict(geo)
            full.update({k: f[k] for k in ("dxu", "dyu", "HTN", "HTE", "tarear")})
            rmm, rtm = mm.copy(), tm.copy()
            rstop = orc.horizontal_remap(case.dom, dt, tt, dep, hd, full, got.io["uvel"], got.io["vvel"], rmm, rtm)
            if rstop[0] == 0:
                assert all(s == rstop for s in stops), (stops, rstop)
                assert np.array_equal(nmm, rmm), "mm differs"
                assert np.array_equal(ntm, rtm) and np.array_equal(np.signbit(ntm), np.signbit(rtm)), "tm differs"
            else:
                # the reference aborts; on several ranks only the rank(s) that hold the offending cell report it
                n_stop += 1
                assert any(s[0] == rstop[0] for s in stops), (stops, rstop)
                if nr == 1:
                    assert stops[0] == rstop, (stops, rstop)
            n_ok += 1
        except Exception as ex:                               # noqa: BLE001
            n_bad += 1
            print("FAIL", info, "->", type(ex).__name__, str(ex)[:300], flush=True)
print("ok", n_ok, "(of which stops:", n_stop, ") bad", n_bad)
