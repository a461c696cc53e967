#!/usr/bin/env python
"""Random small domains through the host build of the device sources (tests/emu) against the CPU oracle, bit for
bit: tripole / displaced-pole / closed-basin grids of 4..78 x 4..29 cells, wet or dry poles, full / caps / half /
no ice, 1-4 ranks, classic / revised EVP, both tilt modes, a turning angle, the coupler's wind stress (strax/stray).

    python tools/fuzz_on_emulator.py SEED [tma|sweep] [CASES]
    FUZZ_THREADS=1 python tools/fuzz_on_emulator.py 5 tma 40      # ranks on threads: the persistent loop kernel with its
                                                                  # tile counters and ghost mailbox between ranks
"""
import sys, time, os, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from cice5_b200 import fixtures as fx
from oracle import oracle as orc
from tests.emu import emu
from tests.util import clone_case, compare_cases
from tests.test_gpu_parity import OUT_CHECK
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
kernel = sys.argv[2] if len(sys.argv) > 2 else "tma"
n_ok = n_bad = 0
defines = tuple(d for d in os.environ.get("FUZZ_DEFINES", "").split() if d)
import threading
def on_threads(ranks, dt):
    errs = []
    def run(e):
        try: e.compute(dt)
        except Exception as ex: errs.append(ex)
    th = [threading.Thread(target=run, args=(e,)) for e in ranks]
    [x.start() for x in th]; [x.join() for x in th]
    if errs: raise errs[0]
with emu.patched(kernel=kernel, defines=defines) as evp:
    if os.environ.get("FUZZ_THREADS"): evp.compute_group = on_threads
    for it in range(int(sys.argv[3]) if len(sys.argv) > 3 else 30):
        tripole = rng.random() < 0.6
        nx = int(rng.integers(2, 40)) * 2 if tripole else int(rng.integers(4, 70))
        ny = int(rng.integers(4, 30))
        scen = ["full", "caps"][int(rng.integers(0, 2))]
        ndte = int(rng.integers(1, 6))
        wet = bool(tripole and rng.random() < 0.5)
        try:
            closed = bool(not tripole and rng.random() < 0.35)
            grid = fx.synthetic_tripole(nx, ny, wet_poles=wet) if tripole else (
                fx.synthetic_closed(nx, ny) if closed else fx.synthetic_displaced(nx, ny))
            case = fx.make_case(grid, scen, ndte=ndte, dt=3600.0)
            mode = rng.random()
            if mode < 0.15:
                for k in ("aice", "vice", "vsno", "aice_init"): case.inp[k][...] = 0.0      # no ice at all
            elif mode < 0.3:
                for k in ("aice", "vice", "vsno", "aice_init"): case.inp[k][ny // 2:, :] = 0.0
            kw = {}
            if rng.random() < 0.25:
                case.inp["strax"] = 0.7 * case.inp["strairyT"] + 0.01
                case.inp["stray"] = -1.3 * case.inp["strairxT"]
            if rng.random() < 0.3: kw["revised_evp"] = True
            if rng.random() < 0.3: kw["tilt_mode"] = 1
            if rng.random() < 0.3: kw.update(cosw=0.9, sinw=float(np.sqrt(1 - 0.81)))
            ref = clone_case(case); orc.run_case(ref, **kw)
            ranks = [r for r in (1, 2, 3, 4) if nx % r == 0 and nx // r >= 2]
            nr = int(rng.choice(ranks))
            got = clone_case(case)
            if nr == 1:
                evp.run_case(got, strict=True, **kw)
            else:
                if kw.get("revised_evp"):
                    tm = grid.f["tmask"][1:-1, 1:-1] > 0
                    kw["xmin"] = float(min(grid.f["dxt"][1:-1, 1:-1][tm].min(), grid.f["dyt"][1:-1, 1:-1][tm].min())) if tm.any() else 1.0
                g = evp.EvpGroup(case.dom, case.ndte, case.dt, nr, strict=True, **kw)
                try:
                    g.set_grid(grid.f); g.evp(case.dt, got.inp, got.io, got.out)
                finally:
                    g.close()
            compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
            n_ok += 1
        except Exception as e:
            n_bad += 1
            print("FAIL", kw, dict(tripole=tripole, nx=nx, ny=ny, scen=scen, ndte=ndte, wet=wet), "->", type(e).__name__, str(e)[:300], flush=True)
print("ok", n_ok, "bad", n_bad)
