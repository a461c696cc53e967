"""The device path against the REFERENCE'S OWN TEXT, without the oracle in between.

tests/golden/ref_driver_vectors.npz holds the outputs of the reference's driver `evp`
(source/ice_dyn_evp.F90:68-510) executed mechanically, statement by statement, from the reference
checkout on multi-block tripole / open domains with the reference's own halo code
(tests/golden/make_driver_vectors.py, tests/golden/fortran_exec.py).  Here the CUDA library computes
the same calls through the C ABI (strict build: no FMA contraction) and every inout / out field must
equal the reference-text result BIT FOR BIT on every physical cell, and uvel / vvel / strength on
every ghost cell -- default build, -Dcoupled (tilt from the sea-surface slope), revised EVP, ice on
the tripole's poles (the 3-D halo routine's unaveraged pole pair), on 1 and 2 ranks.

The strength is handed in from the host (the oracle's ice_strength on every cell, itself pinned bit for
bit to the reference's ice_strength text) and masked to the icetmask list on the device
(strength_masked = 1), which is what the in-driver call of the reference does.

The domains are far smaller than the other device tests' (24 x 12).  Validated on hardware since the end of
round 1 (no xfail mark: a failure is a failure)."""
import os

import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.test_oracle_reference_text import DRIVER_CASES, DRIVER_NPZ, HALOED, blocks_from_global, driver_case

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300), pytest.mark.usefixtures("subcycle_kernel")]


@pytest.fixture(scope="module")
def dvec():
    return np.load(DRIVER_NPZ)


def _check(dvec, name, case, layout):
    _bsx, _bsy, nbx, nby = layout
    checked = 0
    for key in dvec.files:
        if not key.startswith(name + "__out__"):
            continue
        n = key.split("__")[2]
        ref = dvec[key]
        mine = case.io[n] if n in case.io else case.out["strength_out" if n == "strength" else n]
        mine = blocks_from_global(mine.astype(np.float64), layout)
        assert np.array_equal(mine[:, 1:-1, 1:-1], ref[:, 1:-1, 1:-1]), (n, "physical cells")
        if n in HALOED or nbx * nby == 1:
            assert np.array_equal(mine, ref), (n, "ghost cells")
        checked += 1
    assert checked == 34


@pytest.mark.parametrize("name", DRIVER_CASES)
def test_device_matches_reference_text_one_rank(dvec, name):
    case, kw, layout = driver_case(dvec, name)
    evp.run_case(case, strict=True, **kw)
    _check(dvec, name, case, layout)


@pytest.mark.parametrize("name", ["tripole_2x2", "tripole_wet_4x1", "open_3x2", "tripole_access_2x1", "tripole_auscom_2x2", "closed_2x2"])
def test_device_matches_reference_text_two_ranks(dvec, name):
    case, kw, layout = driver_case(dvec, name)
    g = evp.EvpGroup(case.dom, case.ndte, case.dt, 2, strict=True, **kw)
    try:
        g.set_grid(case.grid.f)
        g.evp(case.dt, case.inp, case.io, case.out)
    finally:
        g.close()
    _check(dvec, name, case, layout)


@pytest.mark.parametrize("name,nranks", [("gx3_real_1x1", 1), ("gx3_real_2x2", 2), ("tp1_4x2", 1), ("tp1_4x2", 4),
                                         ("tp1_auscom_4x2", 1), ("tp1_auscom_4x2", 2)])
def test_device_matches_reference_digests(name, nranks):
    """The reference's real gx3 grid and the 1-degree tripole: the reference-text outputs are kept as SHA-256
    digests (tests/golden/ref_driver_digests.json), the inputs are regenerated from the fixtures."""
    from tests.test_oracle_reference_text import check_digests, digest_case
    case, kw, layout, d = digest_case(name)
    if nranks == 1:
        evp.run_case(case, strict=True, **kw)
    else:
        g = evp.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True, **kw)
        try:
            g.set_grid(case.grid.f)
            g.evp(case.dt, case.inp, case.io, case.out)
        finally:
            g.close()
    check_digests(case, layout, d)
