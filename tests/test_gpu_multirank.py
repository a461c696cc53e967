"""x-slab decomposition (one block per rank, NVLink-style inbox halos) on ONE device: all ranks are
driven through evp_gpu_compute_group, which interleaves their steps on one stream.  Results must
be BITWISE equal to the single-block oracle (the reference guarantees decomposition independence,
cicedoc.tex:4601-4603)."""
import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.util import clone_case, compare_cases
from tests.test_gpu_parity import OUT_CHECK, _oracle, _small_tripole

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("subcycle_kernel")]


def _run_group(case, nranks, calls=1, **kw):
    got = clone_case(case)
    g = evp.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True, **kw)
    try:
        g.set_grid(case.grid.f)
        for _ in range(calls):
            got.out = fx.zeroed_outputs(case.dom)
            st = g.evp(case.dt, got.inp, got.io, got.out)
    finally:
        g.close()
    return got, st


@pytest.mark.parametrize("nx,ny,nranks", [(96, 40, 2), (96, 40, 4), (96, 40, 8), (124, 33, 4)])
def test_small_tripole_group_bitwise(nx, ny, nranks):
    case = fx.make_case(_small_tripole(nx, ny), "full", ndte=12, dt=3600.0)
    ref = _oracle(case)
    got, _ = _run_group(case, nranks)
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)


@pytest.mark.parametrize("grid,nranks", [("gx3", 2), ("gx3", 5), ("tp1", 2), ("tp1", 8)])
def test_named_grids_group_bitwise(grid, nranks):
    case = fx.make_case(grid, "caps", ndte=120)
    ref = _oracle(case)
    got, st = _run_group(case, nranks)
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    assert sum(s.active_u_cells for s in st) == int(ref.io["iceumask"][1:-1, 1:-1].sum())


def test_group_two_calls_odd_ndte():
    case = fx.make_case(_small_tripole(96, 40), "full", ndte=5, dt=3600.0)
    ref = clone_case(case)
    from oracle import oracle as orc
    for _ in range(2):
        ref.out = fx.zeroed_outputs(case.dom)
        orc.run_case(ref)
    got, _ = _run_group(case, 4, calls=2)
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)


@pytest.mark.parametrize("grid,starts", [
    ((96, 40), [0, 10, 50, 53]),                 # very uneven, a 3-column slab
    ((124, 33), [0, 2, 4, 70, 72, 100]),         # 2-column slabs next to wide ones
    ("tp1", None),                               # balanced by ocean cells, 8 ranks
])
def test_unequal_slabs_group_bitwise(grid, starts):
    """Slabs of different widths (evp_config_t.slab_start: load balancing by ocean cells, as the
    reference's ice_distribution does): same bits as the single block."""
    g = _small_tripole(*grid) if isinstance(grid, tuple) else fx.make_grid(grid)
    case = fx.make_case(g, "full", ndte=12 if isinstance(grid, tuple) else 40, dt=3600.0)
    if starts is None:
        starts = fx.balanced_slab_starts(g.f["tmask"][1:-1, 1:-1] > 0, 8, strip=31)
        assert len(set(np.diff(starts + [g.dom.nx]))) > 1          # really unequal
    ref = _oracle(case)
    got, st = _run_group(case, len(starts), slab_start=starts)
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    assert sum(s.active_u_cells for s in st) == int(ref.io["iceumask"][1:-1, 1:-1].sum())
