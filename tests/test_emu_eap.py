"""EAP (kdyn = 2) on the host build of the product's own device sources (tests/emu), bit for bit against the oracle.

The oracle's EAP part is pinned to the reference text by tests/test_oracle_eap_reference_text.py; here the device
kernels of cice5_b200/csrc/evp_eap.cuh (eap_subcycle_kernel, eap_stepa_kernel, eap_reset_kernel) and the EAP branch
of the program builder run through the C ABI exactly as in the `-m gpu` tests (tests/test_gpu_zs_eap.py)."""
import numpy as np
import pytest

from cice5_b200 import fixtures as fx
from oracle import oracle as orc
from tests.eap_util import EAP_FIELDS, clone_eap, compare_eap, eap_state
from tests.emu import emu
from tests.util import clone_case

pytestmark = [pytest.mark.timeout(900)]


@pytest.fixture(scope="module")
def dev():
    with emu.patched(kernel="tma") as evp:       # kdyn = 2 selects its own kernel whatever EVP_B200_KERNEL says
        yield evp


@pytest.fixture(scope="module")
def tables():
    return orc.eap_tables()


@pytest.mark.parametrize("grid,scenario,ndte,fresh", [
    ("gx3", "full", 12, False),          # stepa on subcycles 1 and 11
    ("gx3", "caps", 3, True),            # init_eap's isotropic state: atan2(0, 0)
    ("tripole", "full", 11, False),      # stepa on the last subcycle
])
def test_one_rank_bitwise(dev, tables, grid, scenario, ndte, fresh):
    g = fx.synthetic_tripole(96, 40, wet_poles=True) if grid == "tripole" else fx.make_grid(grid)
    case = fx.make_case(g, scenario, ndte=ndte, dt=3600.0)
    e0 = eap_state(case.dom, fresh=fresh)
    ref, ref_e = clone_case(case), clone_eap(e0)
    orc.run_case_eap(ref, ref_e, tables)
    got, got_e = clone_case(case), clone_eap(e0)
    st = dev.run_case_eap(got, got_e, tables, strict=True)
    compare_eap(got, ref, got_e, ref_e, case.dom)
    assert st.active_t_cells > 0
    assert np.count_nonzero(ref_e["yieldstress11"]) > 100 and not ref.out["rdg_shear"].any()
    assert not np.array_equal(ref_e["a11_1"], e0["a11_1"])
    # EAP is not EVP under another name
    evp_ref = clone_case(case)
    orc.run_case(evp_ref)
    assert not np.array_equal(evp_ref.io["uvel"], ref.io["uvel"])


def test_outside_the_regime_nearest_table_entry(dev, tables):
    """ndte = 3 with dt = 3600 s: a subcycle step of the order of 1 / kfrac lets stepa leave the set of valid structure
    tensors (an eigenvalue above 1), where the reference would index past its tables (it does not check).  Device and
    oracle take the nearest entry: no fault, and still the same bits."""
    case = fx.make_case(fx.make_grid("gx3"), "full", ndte=3, dt=3600.0)
    e0 = eap_state(case.dom)
    ref, ref_e = clone_case(case), clone_eap(e0)
    orc.run_case_eap(ref, ref_e, tables)
    lam = max(float((0.5 + np.sqrt((ref_e[f"a11_{c}"] - 0.5) ** 2 + ref_e[f"a12_{c}"] ** 2)).max()) for c in (1, 2, 3, 4))
    assert lam > 1.0                                  # the case does leave the valid set
    got, got_e = clone_case(case), clone_eap(e0)
    dev.run_case_eap(got, got_e, tables, strict=True)
    compare_eap(got, ref, got_e, ref_e, case.dom)


@pytest.mark.parametrize("nranks,starts", [(2, None), (3, [0, 20, 60])])
def test_ranks_two_calls_bitwise(dev, tables, nranks, starts):
    """x-slabs, state (velocities, stresses, structure tensor) carried on the device from one call to the next."""
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    case = fx.make_case(grid, "full", ndte=11, dt=3600.0)
    e0 = eap_state(case.dom)
    ref, got, ref_e, got_e = clone_case(case), clone_case(case), clone_eap(e0), clone_eap(e0)
    g = dev.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True, slab_start=starts, kdyn=2)
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


def test_cuda_graph_and_call_order(dev, tables):
    grid = fx.synthetic_tripole(48, 24)
    case = fx.make_case(grid, "full", ndte=11, dt=3600.0)
    e0 = eap_state(case.dom)
    ref, got, ref_e, got_e = clone_case(case), clone_case(case), clone_eap(e0), clone_eap(e0)
    e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True, use_cuda_graph=True, kdyn=2)
    try:
        e.set_grid(grid.f)
        with pytest.raises(dev.EvpError):           # no tables yet: call order violated
            e.evp(case.dt, got.inp, got.io, got.out)
        e.set_eap_tables(tables)
        e.upload_eap(got_e)
        for _ in range(3):
            ref.out, got.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
            orc.run_case_eap(ref, ref_e, tables)
            e.evp(case.dt, got.inp, got.io, got.out)
            e.download_eap(got_e)
            compare_eap(got, ref, got_e, ref_e, case.dom)
    finally:
        e.close()
    # an EVP handle refuses the EAP entry points; EAP with revised EVP is not a configuration of the reference's eap
    e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    try:
        with pytest.raises(dev.EvpError):
            e.set_eap_tables(tables)
    finally:
        e.close()
    with pytest.raises(dev.EvpError):
        dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True, kdyn=2, revised_evp=True)


def test_bench_eap_block_on_the_host_build(dev, monkeypatch):
    """bench.py's `eap` block (the 1-degree call compared with the oracle, the bench grid timed) with both grids replaced by
    a small tripole: the block's host logic, its parity verdict and the shape of what it adds to the JSON line."""
    import sys
    import bench
    small = fx.synthetic_tripole(48, 24, wet_poles=True)
    monkeypatch.setattr(fx, "make_grid", lambda name: small)
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    args = bench.parse()
    blk = bench.eap_block(args, 0)
    assert set(blk) == {"tp1", "tp01"}
    one = blk["tp1"]
    assert one["parity"]["bitwise"] is True and one["parity"]["fields"] == 55 and not one["parity"]["differing"]
    assert one["ndte"] == 120 and one["value"] > 0 and one["cpu_port"]["value"] > 0
    assert blk["tp01"]["ndte"] == 11 and blk["tp01"]["loop_kernel_launches"] == 2 * 11 + 2 + 11   # yield + sweep, stepa x 2, fold
    assert "kdyn=2 (EAP)" in one["workload"]
