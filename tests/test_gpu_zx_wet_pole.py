"""The velocity fold next to the tripole's pole, on the device against the CPU oracle.

reference: the velocity pair fld2 is updated through ice_HaloUpdate3DR8 (source/ice_dyn_evp.F90:314,
:397), whose NE-corner fold of the MPI build starts at column nxGlobal/2 - lenUsedTripoleBuf + 1 = 2
(mpi/ice_boundary.F90:2817, :1067-1071): the column pair (1, nx-1) of the top row is mirrored with its
sign but not averaged.  Found by executing the reference text (tests/golden/make_halo_vectors.py) and
reproduced in oracle/evp_oracle.c (halo_r8_ex) and csrc/evp_kernels.cuh (fold_necorner_column,
kFirstPair3D).  On every real grid, and on the other fixtures of this suite, those U points are land
and hold 0 either way; `synthetic_tripole(..., wet_poles=True)` puts ice on them.

The device change is two comparisons in fold_necorner_column.  Validated on hardware since the end of round 1
(no xfail mark: a failure is a failure)."""
import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.util import clone_case, compare_cases, phys
from tests.test_gpu_parity import OUT_CHECK

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300), pytest.mark.usefixtures("subcycle_kernel")]


def _case():
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    return grid, fx.make_case(grid, "full", ndte=12, dt=3600.0)


def test_wet_pole_one_rank_bitwise():
    from oracle import oracle as orc
    grid, case = _case()
    nx, ny = case.dom.nx, case.dom.ny
    ref = clone_case(case)
    orc.run_case(ref)
    got = clone_case(case)
    evp.run_case(got, strict=True)
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    assert np.array_equal(phys(got.io["iceumask"]), phys(ref.io["iceumask"]))
    top = got.io["uvel"][ny, 1:nx + 1]
    assert top[0] != 0.0 and top[0] != -top[nx - 2]            # the pole pair really is exercised


@pytest.mark.parametrize("nranks", [2, 4])
def test_wet_pole_ranks_two_calls_bitwise(nranks):
    from oracle import oracle as orc
    grid, case = _case()
    ref = clone_case(case)
    got = clone_case(case)
    g = evp.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True)
    try:
        g.set_grid(grid.f)
        for _ in range(2):
            ref.out = fx.zeroed_outputs(case.dom)
            got.out = fx.zeroed_outputs(case.dom)
            orc.run_case(ref)
            g.evp(case.dt, got.inp, got.io, got.out)
            compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    finally:
        g.close()
