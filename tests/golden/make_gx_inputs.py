"""Extract the reference's shipped gx3 grid / kmt and gx1 kmt binaries into npz inputs.

Run in the build container only (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_gx_inputs.py
These are INPUT data (grid lengths in cm, latitudes in radians, land mask), not reference
source and not reference outputs.  Record layout follows popgrid (source/ice_grid.F90:536-620):
7 direct-access records of nx*ny big-endian float64: ULAT, ULON, HTN, HTE, HUS, HUW, ANGLE.
"""
import numpy as np

REF = "/root/reference/input_templates"
OUT = "cice5_b200/data"


def main():
    g = np.fromfile(f"{REF}/gx3/global_gx3.grid", dtype=">f8").reshape(7, 116, 100)
    kmt3 = np.fromfile(f"{REF}/gx3/global_gx3.kmt", dtype=">i4").reshape(116, 100)
    np.savez_compressed(f"{OUT}/gx3_grid.npz", ULAT=g[0].astype("f8"), HTN_cm=g[2].astype("f8"),
                        HTE_cm=g[3].astype("f8"), kmt=kmt3.astype("i4"))
    kmt1 = np.fromfile(f"{REF}/gx1/global_gx1.kmt", dtype=">i4").reshape(384, 320)
    np.savez_compressed(f"{OUT}/gx1_kmt.npz", kmt=kmt1.astype("i4"))


if __name__ == "__main__":
    main()
