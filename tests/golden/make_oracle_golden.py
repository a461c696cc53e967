"""Regression vectors produced by the CPU oracle itself (NOT by the reference: the Fortran cannot
be built here).  They pin the oracle against accidental change and give the GPU
tests a fixture that needs neither /root/reference nor a second oracle run.
    python tests/golden/make_oracle_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from cice5_b200 import fixtures as fx  # noqa: E402
from oracle import oracle as orc  # noqa: E402

FIELDS_IO = ("uvel", "vvel", "stressp_1", "stressm_2", "stress12_3", "iceumask")
FIELDS_OUT = ("strocnxT", "strocnyT", "divu", "shear", "strintx", "prs_sig")


def main():
    for name, grid, scen in (("gx3_full", "gx3", "full"), ("tp62x40_full", fx.synthetic_tripole(62, 40), "full")):
        case = fx.make_case(grid, scen, ndte=120, dt=3600.0)
        orc.run_case(case)
        d = {f"io_{k}": case.io[k] for k in FIELDS_IO}
        d.update({f"out_{k}": case.out[k] for k in FIELDS_OUT})
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"{name}_ndte120_oracle.npz"), **d)
        print(name, {k: float(np.abs(v).max()) for k, v in d.items()})


if __name__ == "__main__":
    main()
