"""bench.py's measurement function on the host build of the device sources (tests/emu): the parity leg must
say `bitwise: true` for a correct library and `false` when a result is corrupted, the value / e2e checks must
carry umax and the finite flag, and the JSON must hold the keys the driver reads.  No GPU, tiny grid."""
import types

import numpy as np
import pytest

import bench
from cice5_b200 import fixtures as fx
from tests.emu import emu


class _Ctx:
    rank, world, local_rank, dist = 0, 1, 0, None

    def barrier(self):
        pass

    def allmax(self, x):
        return x

    allmin = allsum = allmax

    def gather(self, o):
        return [o]

    def bcast(self, o):
        return o


@pytest.fixture()
def tiny(monkeypatch):
    monkeypatch.setitem(fx.GRIDS, "tiny", lambda: fx.synthetic_tripole(96, 40, "tiny", wet_poles=True))
    monkeypatch.setitem(fx.GRID_DT, "tiny", 3600.0)
    return types.SimpleNamespace(ndte=5, fma=False, no_graph=False, no_compact_metrics=False, equal_slabs=False,
                                 grid="tiny", scenario="full", no_transport=False)


def test_measure_reports_parity_value_checks_and_e2e(tiny):
    with emu.patched(kernel="tma"):
        r = bench.measure(_Ctx(), tiny, "tiny", "full", steps=2, warmup=1, do_e2e=True, do_parity=True)
    assert r["parity"]["bitwise"] is True and r["parity"]["max_rel"] == 0.0 and r["parity"]["fields"] == 36
    assert r["parity"]["nranks"] == 1 and r["parity"]["ndte"] == 5 and r["parity"]["umax"] > 0
    assert r["value_check"]["all_finite"] and r["value_check"]["umax"] > 0
    assert r["e2e"]["all_finite"] and r["e2e"]["umax"] > 0 and r["e2e"]["h2d_bytes_per_step"] > 0
    assert r["loop_kernel_launches_per_step"] == 3 and r["value"] > 0 and r["roofline"]["frac"] > 0
    for k in ("metric", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "dtype",
              "config", "roofline", "e2e", "gpu_launches"):
        assert k in r
    t = r["transport"]                                   # the transport-stage block of the N = 1 bench line
    assert "error" not in t and t["ncat"] == 5 and t["ntrace"] == 11
    for k in ("departure_points", "construct_fields"):
        assert t[k]["kernel_ms"] > 0 and t[k]["achieved_GBps"] > 0 and t[k]["alg_bytes"] > 0
    assert t["construct_fields"]["halo_updates"] == 183
    assert t["horizontal_remap"]["remap_ms"] > 0 and t["horizontal_remap"]["stop"] == [0, 0, 0, 0]
    assert t["horizontal_remap"]["max_abs_area_change"] > 0 and t["horizontal_remap"]["oracle_check"] is None


def test_parity_leg_detects_a_wrong_result(tiny, monkeypatch):
    """One bit flipped in what the oracle returns must turn the verdict to false with a non-zero max_rel."""
    from oracle import oracle as orc
    real = orc.run_case

    def corrupted(case, **kw):
        st = real(case, **kw)
        v = case.io["uvel"]
        j, i = np.unravel_index(np.abs(v).argmax(), v.shape)
        v[j, i] = np.nextafter(v[j, i], np.inf)
        return st

    monkeypatch.setattr(orc, "run_case", corrupted)
    with emu.patched(kernel="tma"):
        r = bench.measure(_Ctx(), tiny, "tiny", "full", steps=1, warmup=0, do_e2e=False, do_parity=True)
    assert r["parity"]["bitwise"] is False and (0, "uvel") in [tuple(m) for m in r["parity"]["mismatches"]]
    assert 0.0 < r["parity"]["max_rel"] < 1e-12


def test_transport_block_oracle_leg(tiny):
    """The transport block with the oracle port beside it (what bench.py attaches to the 1-degree configuration): parity
    of every cell and a CPU time."""
    from cice5_b200 import fixtures
    with emu.patched(kernel="tma") as evp:
        grid = fixtures.make_grid("tiny")
        case = fixtures.make_case(grid, "full", ndte=5, dt=3600.0)
        e = evp.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
        try:
            e.set_grid(grid.f)
            e.evp(case.dt, case.inp, case.io, case.out)
            t = bench.transport_block(e, tiny, grid, case, 40 * case.dt, steps=1, warmup=1, cpu=True)
        finally:
            e.close()
    c = t["horizontal_remap"]["oracle_check"]
    assert c["bitwise"] is True and c["cpu_port_ms"] > 0 and t["horizontal_remap"]["max_abs_area_change"] > 1e-4
