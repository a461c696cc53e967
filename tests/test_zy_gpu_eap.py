"""EAP (kdyn = 2) on the device against the CPU oracle, bit for bit.

reference: eap (source/ice_dyn_eap.F90:66-486), stress_eap (:1052-1467), update_stress_rdg (:1474-1658), stepa
(:1664-1787), calc_ffrac (:1795-1864).  The oracle's EAP part is pinned to the reference text
(tests/test_oracle_eap_reference_text.py); the same kernels run on the host build in tests/test_emu_eap.py.
Collected after the other device suites (and before tests/test_zz_no_alloc.py) so that under `pytest -x` a defect
in this newest row cannot hide the validated ones; each test runs in a child process (`isolated`)."""
import numpy as np
import pytest

from cice5_b200 import evp, fixtures as fx
from tests.eap_util import clone_eap, compare_eap, eap_state
from tests.util import clone_case

pytestmark = [pytest.mark.gpu, pytest.mark.isolated, pytest.mark.timeout(300)]


@pytest.mark.parametrize("grid,scenario,ndte", [("gx3", "full", 120), ("tp1", "caps", 21)])
def test_eap_one_call_bitwise(grid, scenario, ndte):
    from oracle import oracle as orc
    tables = orc.eap_tables()
    case = fx.make_case(grid, scenario, ndte=ndte)
    e0 = eap_state(case.dom)
    ref, ref_e = clone_case(case), clone_eap(e0)
    orc.run_case_eap(ref, ref_e, tables)
    got, got_e = clone_case(case), clone_eap(e0)
    st = evp.run_case_eap(got, got_e, tables, strict=True)
    compare_eap(got, ref, got_e, ref_e, case.dom)
    assert st.active_t_cells > 0 and np.count_nonzero(ref_e["yieldstress11"]) > 100
    print(f"EAP {grid} {scenario} ndte={ndte}: loop {st.loop_ms:.3f} ms = {1e3 * st.loop_ms / ndte:.1f} us per subcycle, "
          f"{st.active_t_cells} T cells")


def test_eap_tenth_degree_bitwise():
    """The bench grid (3600 x 2700, 7.6 M active T cells), ndte = 11: stepa on subcycles 1 and 11.  The oracle needs about
    13 s for this call on 16 host threads (profiles/r02_eap.md: first run through tools/profile_eap.py, calls 42 / 43)."""
    from oracle import oracle as orc
    tables = orc.eap_tables()
    case = fx.make_case("tp01", "full", ndte=11)
    e0 = eap_state(case.dom)
    ref, ref_e = clone_case(case), clone_eap(e0)
    orc.run_case_eap(ref, ref_e, tables)
    got, got_e = clone_case(case), clone_eap(e0)
    st = evp.run_case_eap(got, got_e, tables, strict=True)
    compare_eap(got, ref, got_e, ref_e, case.dom)
    assert st.active_t_cells > 7_000_000
    print(f"EAP tp01 full ndte=11: loop {st.loop_ms:.3f} ms = {1e3 * st.loop_ms / 11:.1f} us per subcycle")


def test_eap_two_calls_three_ranks_and_graph_bitwise():
    from oracle import oracle as orc
    tables = orc.eap_tables()
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    case = fx.make_case(grid, "full", ndte=11, dt=3600.0)
    e0 = eap_state(case.dom)
    ref, got, ref_e, got_e = clone_case(case), clone_case(case), clone_eap(e0), clone_eap(e0)
    g = evp.EvpGroup(case.dom, case.ndte, case.dt, 3, strict=True, slab_start=[0, 20, 60], kdyn=2)
    try:
        g.set_grid(grid.f)
        g.set_eap_tables(tables)
        g.upload_eap(got_e)
        for _ in range(2):
            ref.out, got.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
            orc.run_case_eap(ref, ref_e, tables)
            g.evp(case.dt, got.inp, got.io, got.out)
            g.download_eap(got_e)
            compare_eap(got, ref, got_e, ref_e, case.dom)
    finally:
        g.close()
    # one rank, the loop as a CUDA graph, isotropic start
    e0 = eap_state(case.dom, fresh=True)
    ref, got, ref_e, got_e = clone_case(case), clone_case(case), clone_eap(e0), clone_eap(e0)
    e = evp.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True, use_cuda_graph=True, kdyn=2)
    try:
        e.set_grid(grid.f)
        e.set_eap_tables(tables)
        for _ in range(2):
            ref.out, got.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
            orc.run_case_eap(ref, ref_e, tables)
            e.evp(case.dt, got.inp, got.io, got.out)
            e.download_eap(got_e)
            compare_eap(got, ref, got_e, ref_e, case.dom)
    finally:
        e.close()
