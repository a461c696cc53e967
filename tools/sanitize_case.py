#!/usr/bin/env python
"""One small evp(dt) through the C ABI for compute-sanitizer runs (memcheck / racecheck / synccheck):
   compute-sanitizer --tool memcheck python tools/sanitize_case.py gx3 tma 4
Checks the result against the oracle so that a tool-induced slowdown cannot hide a wrong answer."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
grid, kernel, ndte = sys.argv[1], sys.argv[2], int(sys.argv[3])
nranks = int(sys.argv[4]) if len(sys.argv) > 4 else 1
os.environ["EVP_B200_KERNEL"] = kernel
import numpy as np
from cice5_b200 import evp, fixtures as fx
from oracle import oracle as orc
from tests.util import clone_case, compare_cases, STATE

if grid == "tp96x40":
    g = fx.synthetic_tripole(96, 40, "tp96x40", wet_poles=True)
else:
    g = fx.make_grid(grid)
case = fx.make_case(g, "caps" if grid.startswith("tp") else "full", ndte=ndte)
ref = clone_case(case); orc.run_case(ref)
got = clone_case(case)
if nranks == 1:
    st = evp.run_case(got, strict=True, device=0)
else:
    grp = evp.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True)
    try:
        grp.set_grid(g.f)
        grp.evp(case.dt, got.inp, got.io, got.out)
        st = grp.ranks[0].stats()
    finally:
        grp.close()
compare_cases(got, ref, fields_io=STATE, fields_out=("divu", "shear", "strocnxT", "strocnyT"), bitwise=True)
print(f"sanitize_case {grid} {kernel} ndte={ndte} nranks={nranks}: bitwise == oracle, "
      f"{st.kernel_launches} launches, umax {np.abs(got.io['uvel']).max():.4f}")
