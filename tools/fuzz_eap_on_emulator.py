#!/usr/bin/env python
"""EAP (kdyn = 2) on random small domains through the host build of the device sources (tests/emu) against the CPU oracle,
bit for bit on every field of the call and every EAP field: tripole / displaced-pole / closed-basin grids of 4..78 x 4..29
cells, wet or dry poles, full / caps / half / no ice, 1-4 ranks, ndte 1..23 (stepa on subcycles 1, 11, 21), a turning
angle, both tilt modes, a fresh (isotropic) or a seeded structure tensor, one or two calls, CUDA graph on or off.

    python tools/fuzz_eap_on_emulator.py SEED [CASES]
"""
import os
import sys
import time
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from cice5_b200 import fixtures as fx  # noqa: E402
from oracle import oracle as orc  # noqa: E402
from tests.eap_util import clone_eap, compare_eap, eap_state  # noqa: E402
from tests.emu import emu  # noqa: E402
from tests.util import clone_case  # noqa: E402

rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
cases = int(sys.argv[2]) if len(sys.argv) > 2 else 30
tables = fx.eap_tables()
n_ok = n_bad = 0
t0 = time.time()
with emu.patched() as evp:
    for it in range(cases):
        tripole = rng.random() < 0.6
        nx = int(rng.integers(2, 40)) * 2 if tripole else int(rng.integers(4, 70))
        ny = int(rng.integers(4, 30))
        scen = ["full", "caps"][int(rng.integers(0, 2))]
        ndte = int(rng.choice([1, 2, 5, 11, 12, 21, 23]))
        wet = bool(tripole and rng.random() < 0.5)
        desc = ""
        try:
            closed = bool(not tripole and rng.random() < 0.35)
            grid = fx.synthetic_tripole(nx, ny, wet_poles=wet) if tripole else (
                fx.synthetic_closed(nx, ny) if closed else fx.synthetic_displaced(nx, ny))
            # a subcycle step inside the model's regime (dte << 1 / kfrac): dt = 120 s per subcycle
            case = fx.make_case(grid, scen, ndte=ndte, dt=120.0 * ndte)
            mode = rng.random()
            if mode < 0.15:
                for k in ("aice", "vice", "vsno", "aice_init"):
                    case.inp[k][...] = 0.0
            elif mode < 0.3:
                for k in ("aice", "vice", "vsno", "aice_init"):
                    case.inp[k][ny // 2:, :] = 0.0
            kw = {}
            if rng.random() < 0.3:
                kw["tilt_mode"] = 1
            if rng.random() < 0.3:
                kw.update(cosw=0.9, sinw=float(np.sqrt(1 - 0.81)))
            fresh = bool(rng.random() < 0.3)
            calls = int(rng.integers(1, 3))
            graph = bool(rng.random() < 0.5)
            ranks = [r for r in (1, 2, 3, 4) if nx % r == 0 and nx // r >= 2]
            nr = int(rng.choice(ranks))
            desc = (f"{grid.dom.name} {nx}x{ny} {scen} ndte={ndte} ranks={nr} calls={calls} graph={graph} fresh={fresh} "
                    f"mode={mode:.2f} {kw}")
            e0 = eap_state(case.dom, seed=int(rng.integers(0, 1 << 30)), fresh=fresh)
            ref, got, ref_e, got_e = clone_case(case), clone_case(case), clone_eap(e0), clone_eap(e0)
            g = evp.EvpGroup(case.dom, ndte, case.dt, nr, strict=True, kdyn=2, use_cuda_graph=graph, **kw)
            try:
                g.set_grid(grid.f)
                g.set_eap_tables(tables)
                g.upload_eap(got_e)
                for _ in range(calls):
                    ref.out, got.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
                    orc.run_case_eap(ref, ref_e, tables, **kw)
                    g.evp(case.dt, got.inp, got.io, got.out)
                    g.download_eap(got_e)
                    compare_eap(got, ref, got_e, ref_e, case.dom)
            finally:
                g.close()
            n_ok += 1
        except Exception as ex:      # noqa: BLE001
            n_bad += 1
            print("FAIL", desc, "::", str(ex)[:300], flush=True)
            traceback.print_exc(limit=2)
print(f"EAP fuzz: {n_ok} ok, {n_bad} failed of {cases} in {time.time() - t0:.0f} s", flush=True)
sys.exit(1 if n_bad else 0)
