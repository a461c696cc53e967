"""The product's OWN device and ABI sources, compiled for the host and run without a GPU.

tests/emu/ rewrites nothing but the launch syntax, the CUDA headers and three PTX instructions of
cice5_b200/csrc and builds the result with g++ against a small stand-in of the CUDA runtime (streams are
threads, threads of a block are fibers, warp collectives are rendezvous; tests/emu/cuda_emu.h).  The
C ABI library that comes out is driven through cice5_b200.evp exactly as the CUDA library is in the
`-m gpu` tests, and must equal the CPU oracle -- and the vectors made from the reference's own text --
BIT FOR BIT.  What this covers that had no hardware run yet: revised EVP, the device ice_strength,
principal_stress, the velocity fold's unaveraged pole pair, the comparison with the reference-text
vectors, tiny domains; and it re-checks on the host what the GPU suite validated: prep / halo / fold /
finish kernels, the program builder, the CUDA-graph path, the multi-rank flag barrier, unequal slabs.
Both subcycle kernels run: the TMA ring (mbarrier and cp.async.bulk.tensor have host stand-ins: a copy is
carried out when issued and completes its bytes on the barrier) and the register sweep.  What it cannot
cover: the memory model and the asynchrony of a real device (a missing fence, a stage released too early
or a race between CTAs does not show here), and speed.

This is test infrastructure.  The product never loads the host build (nothing under cice5_b200/ refers
to tests/emu/) and keeps failing loudly when its CUDA library is missing."""
import os

import numpy as np
import pytest

from cice5_b200 import fixtures as fx
from oracle import oracle as orc
from tests.emu import emu
from tests.util import STATE, clone_case, compare_cases, phys
from tests.test_gpu_parity import OUT_CHECK

pytestmark = [pytest.mark.timeout(600)]


@pytest.fixture(scope="module", params=["tma", "sweep"])
def dev(request):
    """cice5_b200.evp bound to the host build; every test runs with the TMA ring kernel (the product's
    default: producer warp, mbarriers, tensor-map copies, work queue, item list) and with the register sweep."""
    with emu.patched(kernel=request.param) as evp:
        yield evp


def _oracle(case, **kw):
    ref = clone_case(case)
    orc.run_case(ref, **kw)
    return ref


def test_host_build_exports_the_product_abi(dev):
    import ctypes as C
    lib = C.CDLL(emu.build())
    for name in dev.EXPORTS:
        assert hasattr(lib, name), name


@pytest.mark.parametrize("grid,scenario,ndte,kw", [
    ("gx3", "full", 8, {}),
    ("gx3", "caps", 7, dict(cosw=0.9, sinw=float(np.sqrt(1 - 0.81)), tilt_mode=1)),
    ("tripole", "full", 9, {}),
    ("tripole_wet", "full", 6, {}),
])
def test_one_rank_bitwise(dev, grid, scenario, ndte, kw):
    g = {"tripole": lambda: fx.synthetic_tripole(96, 40), "tripole_wet": lambda: fx.synthetic_tripole(96, 40, wet_poles=True)
         }.get(grid, lambda: fx.make_grid(grid))()
    case = fx.make_case(g, scenario, ndte=ndte, dt=3600.0)
    ref = _oracle(case, **kw)
    got = clone_case(case)
    st = dev.run_case(got, strict=True, **kw)
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    assert np.array_equal(phys(got.io["iceumask"]), phys(ref.io["iceumask"]))
    assert st.active_t_cells > 0 and st.kernel_launches > ndte
    if grid == "tripole_wet":                                   # the 3-D halo routine's pole pair
        top = got.io["uvel"][g.dom.ny, 1:g.dom.nx + 1]
        assert top[0] != 0.0 and top[0] != -top[g.dom.nx - 2]


@pytest.mark.parametrize("nranks,starts", [(2, None), (4, None), (3, [0, 20, 60])])
def test_ranks_two_calls_bitwise(dev, nranks, starts):
    """x-slabs with the in-kernel flag barrier (streams are threads here, so the spin-waits are real),
    equal and unequal slabs, state carried from one call to the next, on the wet-pole tripole."""
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    case = fx.make_case(grid, "full", ndte=5, dt=3600.0)
    ref, got = clone_case(case), clone_case(case)
    g = dev.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True, slab_start=starts)
    try:
        g.set_grid(grid.f)
        for _ in range(2):
            ref.out, got.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
            orc.run_case(ref)
            g.evp(case.dt, got.inp, got.io, got.out)
            compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    finally:
        g.close()


def test_cuda_graph_path_and_odd_ndte(dev):
    """The ndte loop captured once per starting ping-pong parity and replayed: three calls with an odd ndte
    (the state alternates between the two copies of the planes), graph on and off."""
    grid = fx.synthetic_tripole(48, 24)
    case = fx.make_case(grid, "full", ndte=5, dt=3600.0)
    for graph in (True, False):
        ref, got = clone_case(case), clone_case(case)
        e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True, use_cuda_graph=graph)
        try:
            e.set_grid(grid.f)
            for _ in range(3):
                ref.out, got.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
                orc.run_case(ref)
                e.evp(case.dt, got.inp, got.io, got.out)
                compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
        finally:
            e.close()


def _xmin(grid):
    tm = grid.f["tmask"][1:-1, 1:-1] > 0
    return float(min(grid.f["dxt"][1:-1, 1:-1][tm].min(), grid.f["dyt"][1:-1, 1:-1][tm].min()))


def test_revised_evp_bitwise(dev):
    """REVISED instantiations of the subcycle kernel, prep2's stress reset, constants from xmin."""
    grid = fx.make_grid("gx3")
    case = fx.make_case(grid, "full", ndte=6)
    ref = _oracle(case, revised_evp=True)
    got = clone_case(case)
    dev.run_case(got, strict=True, revised_evp=True)
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    classic = _oracle(case)
    assert not np.array_equal(phys(ref.io["uvel"]), phys(classic.io["uvel"]))


def test_revised_evp_two_ranks_two_calls_bitwise(dev):
    grid = fx.synthetic_tripole(96, 40)
    case = fx.make_case(grid, "full", ndte=5, dt=3600.0)
    ref, got = clone_case(case), clone_case(case)
    g = dev.EvpGroup(case.dom, case.ndte, case.dt, 2, strict=True, revised_evp=True, xmin=_xmin(grid))
    try:
        g.set_grid(grid.f)
        for _ in range(2):
            ref.out, got.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
            orc.run_case(ref, revised_evp=True)
            g.evp(case.dt, got.inp, got.io, got.out)
            compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    finally:
        g.close()


@pytest.mark.parametrize("modes", [dict(krdg_partic=1, krdg_redist=1), dict(krdg_partic=0, krdg_redist=0),
                                   dict(krdg_partic=0, krdg_redist=1), dict(krdg_partic=1, krdg_redist=0),
                                   dict(kstrength=0)])
def test_device_ice_strength_bitwise(dev, modes):
    """ice_strength_kernel from the uploaded thickness distribution.  Both sides evaluate the correctly rounded
    evp_exp (csrc/evp_exp.h), so even the exponential participation function must agree bit for bit -- here
    and on a GPU."""
    grid = fx.make_grid("gx3")
    case = fx.make_case(grid, "caps", ndte=4)
    itd = fx.ice_itd(case.inp)
    ref = clone_case(case)
    ref.inp["strength"] = orc.ice_strength(grid.dom, case.inp["aice"], case.inp["vice"], itd["aice0"], itd["aicen"],
                                           itd["vicen"], cr_exp=True, **modes)
    masked = 0 if modes.get("kstrength", 1) == 0 else 1
    orc.run_case(ref, strength_masked=masked)
    got = clone_case(case)
    got.inp.pop("strength")
    e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    try:
        e.set_grid(grid.f)
        e.upload_itd(itd["aicen"], itd["vicen"], itd["aice0"], **modes)
        e.evp(case.dt, got.inp, got.io, got.out)
    finally:
        e.close()
    assert (phys(ref.out["strength_out"]) > 0).any()
    assert np.array_equal(phys(got.out["strength_out"]), phys(ref.out["strength_out"]))
    compare_cases(got, ref, fields_io=STATE, fields_out=("divu", "shear", "strocnxT"), bitwise=True)


def test_principal_stress_bitwise(dev):
    grid = fx.make_grid("gx3")
    case = fx.make_case(grid, "full", ndte=6)
    got = clone_case(case)
    e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    try:
        e.set_grid(grid.f)
        e.evp(case.dt, got.inp, got.io, got.out)
        s1, s2 = e.principal_stress()
    finally:
        e.close()
    r1, r2 = orc.principal_stress(got.io["stressp_1"], got.io["stressm_1"], got.io["stress12_1"], got.out["prs_sig"])
    assert np.array_equal(s1, r1) and np.array_equal(s2, r2)
    assert (s1 != 1.0e30).any() and (s1 == 1.0e30).any()


def test_resident_state_upload_compute_download(dev):
    """upload once, compute twice, download: equals two evp calls of the oracle with the same forcing."""
    grid = fx.synthetic_tripole(48, 24)
    case = fx.make_case(grid, "full", ndte=4, dt=3600.0)
    ref = clone_case(case)
    orc.run_case(ref)
    ref.out = fx.zeroed_outputs(case.dom)
    orc.run_case(ref)
    got = clone_case(case)
    e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    try:
        e.set_grid(grid.f)
        e.upload(got.inp, got.io)
        e.compute(case.dt)
        e.compute(case.dt)
        e.download(got.io, got.out)
    finally:
        e.close()
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)


# ---- the device sources against the reference's own text, no oracle in between ---------------------------
from tests.test_oracle_reference_text import DRIVER_CASES, DRIVER_NPZ, HALOED, blocks_from_global, driver_case  # noqa: E402


@pytest.fixture(scope="module")
def dvec():
    return np.load(DRIVER_NPZ)


def _check_reference_text(dvec, name, case, layout):
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
def test_device_sources_match_reference_text(dev, dvec, name):
    case, kw, layout = driver_case(dvec, name)
    dev.run_case(case, strict=True, **kw)
    _check_reference_text(dvec, name, case, layout)


@pytest.mark.parametrize("name", ["tripole_2x2", "tripole_wet_4x1", "open_3x2", "tripole_access_2x1", "tripole_auscom_2x2", "closed_2x2"])
def test_device_sources_match_reference_text_two_ranks(dev, dvec, name):
    case, kw, layout = driver_case(dvec, name)
    g = dev.EvpGroup(case.dom, case.ndte, case.dt, 2, strict=True, **kw)
    try:
        g.set_grid(case.grid.f)
        g.evp(case.dt, case.inp, case.io, case.out)
    finally:
        g.close()
    _check_reference_text(dvec, name, case, layout)


def test_pinned_host_arrays(dev):
    """evp_gpu_pin_host_arrays: page-locks the caller's arrays once (twice is harmless), results unchanged,
    registrations released with the handle."""
    grid = fx.synthetic_tripole(40, 20)
    case = fx.make_case(grid, "full", ndte=3, dt=3600.0)
    ref, got = clone_case(case), clone_case(case)
    orc.run_case(ref)
    e = dev.EvpGpu.for_domain(got.dom, got.ndte, got.dt, strict=True)
    try:
        e.set_grid(grid.f)
        e.pin_host(got.inp, got.io, got.out)
        e.pin_host(got.inp, got.io, None)
        e.evp(got.dt, got.inp, got.io, got.out)
    finally:
        e.close()
    compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)


def test_strax_without_stray_is_rejected(dev, dvec):
    """evp_in_t.strax / stray (ACCESS, calc_strair = .false.): both or neither."""
    case, kw, _ = driver_case(dvec, "tripole_access_2x1")
    case.inp["stray"] = None
    with pytest.raises(dev.EvpError):
        dev.run_case(case, strict=True, **kw)


def test_product_loader_has_no_host_fallback():
    """cice5_b200 must not know about the host build: no reference to tests/emu anywhere in the package."""
    import os
    pkg = os.path.dirname(fx.__file__)
    for root, _dirs, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".F90")):
                text = open(os.path.join(root, f), errors="replace").read()
                assert "tests/emu" not in text and "libevp_emu" not in text and "cuda_emu" not in text, f


def _compute_on_threads(ranks, dt):
    """One host thread per rank, each on its own stream -- the in-process picture of one process per GPU
    (evp_gpu_compute_group puts all ranks on ONE stream of one device, where a kernel that waits for a peer
    queued behind it can never end; there the fold stays a launch of its own)."""
    import threading
    errors = []

    def run(e):
        try:
            e.compute(dt)
        except Exception as ex:                      # noqa: BLE001
            errors.append(ex)

    threads = [threading.Thread(target=run, args=(e,)) for e in ranks]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    if errors:
        raise errors[0]


def _group_case(evp, monkeypatch, grid, case, nranks, starts=None, threads=True, calls=2, **kw):
    """`calls` evp calls of an EvpGroup against the oracle, bit for bit; ranks on threads (one stream each: the
    persistent loop kernel is the path) or through evp_gpu_compute_group (one stream: a launch per subcycle)."""
    ref, got = clone_case(case), clone_case(case)
    if threads:
        monkeypatch.setattr(evp, "compute_group", _compute_on_threads)
    g = evp.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True, slab_start=starts, **kw)
    try:
        g.set_grid(grid.f)
        for _ in range(calls):
            ref.out, got.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
            orc.run_case(ref, **({"revised_evp": True} if kw.get("revised_evp") else {}))
            stats = g.evp(case.dt, got.inp, got.io, got.out)
            compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
    finally:
        g.close()
    return stats


@pytest.mark.parametrize("nranks,starts,rpc", [(1, None, None), (2, None, None), (4, None, "3"), (3, [0, 20, 60], "5"),
                                               (8, None, "2"), (1, None, "1"), (3, [0, 20, 60], "1")])
def test_loop_kernel_tripole_ranks_on_threads_bitwise(monkeypatch, nranks, starts, rpc):
    """evp_loop_tma_kernel (all ndte subcycles in one launch, per-tile dependency counters, the fold as an item):
    1 to 8 ranks really overlapping (a thread and a stream per rank), equal and unequal slabs, several chunk
    heights, ice on the poles, two calls -- and exactly one loop launch per rank and call."""
    if rpc:
        monkeypatch.setenv("EVP_B200_RPC", rpc)
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    case = fx.make_case(grid, "full", ndte=7, dt=3600.0)
    with emu.patched(kernel="tma") as evp:
        stats = _group_case(evp, monkeypatch, grid, case, nranks, starts)
        # loop kernel + last subcycle + fold; at N > 1 also a signal / wait pair after the loop kernel, the mailbox drain
        # and a signal / wait pair before the fold
        assert all(s.loop_kernel_launches == (3 if nranks == 1 else 8) for s in stats)


@pytest.mark.parametrize("gridname,nranks", [("open", 1), ("open", 3), ("closed", 2), ("caps", 2), ("wide", 2)])
def test_loop_kernel_other_boundaries_on_threads_bitwise(monkeypatch, gridname, nranks):
    """The same without a fold (cyclic / open, closed basin), with tiles that hold no ice and are never run
    (caps scenario: their counters start at all ones), and with several strips per rank (wide: 300 columns)."""
    monkeypatch.setenv("EVP_B200_RPC", "4")
    if gridname == "open":
        grid = fx.synthetic_displaced(60, 30, name="open60")
    elif gridname == "closed":
        grid = fx.synthetic_closed(40, 24, name="closed40")
    elif gridname == "wide":
        grid = fx.synthetic_tripole(300, 18, "wide300", wet_poles=True)
    else:
        grid = fx.synthetic_tripole(96, 40, "caps96")
    case = fx.make_case(grid, "caps" if gridname == "caps" else "full", ndte=6, dt=3600.0)
    with emu.patched(kernel="tma") as evp:
        _group_case(evp, monkeypatch, grid, case, nranks)


def test_loop_kernel_revised_evp_and_loop_switch(monkeypatch):
    """revised EVP through the loop kernel (2 ranks on threads); EVP_B200_LOOP=0 brings back a launch per
    subcycle with the same answers."""
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    case = fx.make_case(grid, "full", ndte=5, dt=3600.0)
    xmin = _xmin(grid)
    with emu.patched(kernel="tma") as evp:
        _group_case(evp, monkeypatch, grid, case, 2, revised_evp=True, xmin=xmin)
        monkeypatch.setenv("EVP_B200_LOOP", "0")
        stats = _group_case(evp, monkeypatch, grid, case, 2)
        assert all(s.loop_kernel_launches == 2 * case.ndte for s in stats)


def test_plain_c_host_example_on_the_host_build(tmp_path):
    """examples/evp_host.c (init -> set_grid -> run -> get_stats -> finalize from C99, no Python in the
    loop) linked against the host build: the numbers the oracle gives for that closed basin."""
    import os
    import re
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = emu.build()
    exe = str(tmp_path / "evp_host_emu")
    r = subprocess.run([cc, "-O2", "-std=c99", "-Wall", "-I", os.path.join(root, "include"),
                        os.path.join(root, "examples", "evp_host.c"), "-L", os.path.dirname(lib), "-levp_emu",
                        f"-Wl,-rpath,{os.path.dirname(lib)}", "-lm", "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"uvel = (\S+)\s+vvel = (\S+)", r.stdout)
    u, v = float(m.group(1)), float(m.group(2))
    assert abs(u - 1.154007e-01) <= 2e-6 * 1.154007e-01 and abs(v + 2.511930e-02) <= 2e-6 * 2.511930e-02
    assert "5828 active T cells" in r.stdout


def test_random_small_domains_bitwise(dev):
    """A dozen random domains per kernel (tools/fuzz_on_emulator.py runs hundreds): sizes down to 4 x 4,
    wet and dry poles, no ice at all, ice in the south only, 1-4 ranks, revised EVP, tilt, turning angle."""
    rng = np.random.default_rng(20261104)
    for _ in range(12):
        tripole = rng.random() < 0.6
        nx = int(rng.integers(2, 30)) * 2 if tripole else int(rng.integers(4, 50))
        ny = int(rng.integers(4, 24))
        grid = fx.synthetic_tripole(nx, ny, wet_poles=bool(rng.random() < 0.5)) if tripole else fx.synthetic_displaced(nx, ny)
        case = fx.make_case(grid, ["full", "caps"][int(rng.integers(0, 2))], ndte=int(rng.integers(1, 5)), dt=3600.0)
        mode = rng.random()
        if mode < 0.15:
            for k in ("aice", "vice", "vsno", "aice_init"):
                case.inp[k][...] = 0.0
        elif mode < 0.3:
            for k in ("aice", "vice", "vsno", "aice_init"):
                case.inp[k][ny // 2:, :] = 0.0
        kw = {}
        if rng.random() < 0.3:
            kw["revised_evp"] = True
        if rng.random() < 0.3:
            kw["tilt_mode"] = 1
        if rng.random() < 0.3:
            kw.update(cosw=0.9, sinw=float(np.sqrt(1 - 0.81)))
        ref = _oracle(case, **kw)
        nr = int(rng.choice([r for r in (1, 2, 3, 4) if nx % r == 0 and nx // r >= 2]))
        got = clone_case(case)
        if nr == 1:
            dev.run_case(got, strict=True, **kw)
        else:
            if kw.get("revised_evp"):
                kw["xmin"] = _xmin(grid)
            g = dev.EvpGroup(case.dom, case.ndte, case.dt, nr, strict=True, **kw)
            try:
                g.set_grid(grid.f)
                g.evp(case.dt, got.inp, got.io, got.out)
            finally:
                g.close()
        compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)


@pytest.mark.parametrize("name", ["gx3_real_1x1", "gx3_real_2x2", "tp1_auscom_4x2"])
def test_device_sources_match_reference_digests_on_the_real_gx3_grid(dev, name):
    """The reference's real gx3 grid file, and the 1-degree tripole in the ACCESS-OM production build: device
    sources == the reference text (SHA-256 of every field)."""
    from tests.test_oracle_reference_text import check_digests, digest_case
    case, kw, layout, d = digest_case(name)
    dev.run_case(case, strict=True, **kw)
    check_digests(case, layout, d)


def _rand_stresses(case, seed=7):
    rng = np.random.default_rng(seed)
    for n in fx.STRESS_FIELDS:
        case.io[n][...] = rng.standard_normal(case.io[n].shape) * 1.0e3      # ghost cells hold garbage too


@pytest.mark.parametrize("nranks", [1, 2, 4])
def test_restart_fixup_bitwise(dev, nranks):
    """evp_gpu_restart_fixup == restartfile's treatment of freshly read stresses (ice_restart_driver.F90:370-394:
    12 x ice_HaloUpdate_stress on a tripole grid; :431-466: south ghost row, west ghost column zeroed), every cell
    of every stress incl. the outer ghost ring, 1 / 2 / 4 ranks; a closed basin too (no halo part)."""
    for grid in (fx.synthetic_tripole(96, 40, wet_poles=True), fx.synthetic_closed(40, 24, name="closed40")):
        case = fx.make_case(grid, "full", ndte=2, dt=3600.0)
        _rand_stresses(case)
        ref, got = clone_case(case), clone_case(case)
        orc.restart_fixup(case.dom, ref.io)
        g = dev.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True)
        try:
            g.set_grid(grid.f)
            g.upload(got.inp, got.io)
            g.restart_fixup()
            g.download(got.io, got.out)
        finally:
            g.close()
        for n in fx.STRESS_FIELDS:
            assert np.array_equal(got.io[n], ref.io[n]), (grid.dom.name, n)
            assert not got.io[n][0, :].any() and not got.io[n][:, 0].any()
        if grid.dom.ns_tripole:
            assert got.io["stressp_1"][-1, 1:-1].any()


def test_resident_protocol_of_evp_gpu_run(dev):
    """evp_gpu_set_resident: the state goes up once, later calls upload only the inputs and bring back only the
    groups of the mask -- the host arrays of the other groups must stay untouched, the downloaded ones must
    equal a full-copy run bit for bit over three calls, and evp_gpu_download afterwards returns everything."""
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    case = fx.make_case(grid, "full", ndte=5, dt=3600.0)
    ref = clone_case(case)
    got = clone_case(case)
    mask = dev.DL_VELOCITY | dev.DL_DEFORM | dev.DL_OCN_T
    minimal = ("uvel", "vvel", "divu", "shear", "rdg_conv", "rdg_shear", "strocnxT", "strocnyT")
    e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    try:
        e.set_grid(grid.f)
        e.set_resident(True, mask)
        for call in range(3):
            orc.run_case(ref)
            got.out = {k: np.full_like(v, -777) for k, v in got.out.items()}
            stale = {k: v.copy() for k, v in got.io.items()}
            if call > 0:                       # resident: whatever the host holds in its state arrays is ignored
                for n in fx.STRESS_FIELDS:
                    got.io[n][...] = 123.0
                stale = {k: v.copy() for k, v in got.io.items()}
            e.evp(case.dt, got.inp, got.io, got.out)
            for n in minimal:
                a = got.io[n] if n in got.io else got.out[n]
                b = ref.io[n] if n in ref.io else ref.out[n]
                assert np.array_equal(phys(a), phys(b)), (call, n)
            for n in fx.STRESS_FIELDS:          # not in the mask: host copy untouched
                assert np.array_equal(got.io[n], stale[n]), n
            assert (got.out["strairx"] == -777).all() and (got.out["prs_sig"] == -777).all()
        e.download(got.io, got.out)              # history / restart step: everything
        compare_cases(got, ref, fields_out=OUT_CHECK, bitwise=True)
        # a restart was read on the host: invalidate, the next call uploads the state again
        e.invalidate_state()
        for k in ref.io:
            got.io[k][...] = case.io[k]
            ref.io[k][...] = case.io[k]
        orc.run_case(ref)
        e.evp(case.dt, got.inp, got.io, got.out)
        assert np.array_equal(phys(got.io["uvel"]), phys(ref.io["uvel"]))
    finally:
        e.close()


# ---- horizontal transport, first stage: departure_points from the resident velocities --------
def _departure_case(dev, grid, scenario, ndte, nranks, midpt, dt_scale=1.0):
    """evp(dt), then departure_points on the device against the oracle's (itself pinned to the reference text), bit for
    bit on every cell of the global block, ghost ring included (1 rank) / on the cells unslab keeps (several ranks)."""
    case = fx.make_case(grid, scenario, ndte=ndte, dt=3600.0)
    got = clone_case(case)
    dt = case.dt * dt_scale
    f = grid.f
    if nranks == 1:
        e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
        try:
            e.set_grid(f)
            e.set_transport_grid(f)
            e.evp(case.dt, got.inp, got.io, got.out)
            dpx, dpy, l_stop, istop, jstop = e.departure_points(dt, midpt)
        finally:
            e.close()
        stops = [(l_stop, istop, jstop)]
    else:
        g = dev.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True)
        try:
            g.set_grid(f)
            g.set_transport_grid(f)
            g.evp(case.dt, got.inp, got.io, got.out)
            dpx, dpy, stops = g.departure_points(dt, midpt)
        finally:
            g.close()
    rx, ry, r_stop, ri, rj = orc.departure_points(case.dom, dt, got.io["uvel"], got.io["vvel"], f["dxu"], f["dyu"], f["HTN"],
                                                  f["HTE"], l_dp_midpt=midpt)
    assert np.abs(got.io["uvel"]).max() > 1e-3
    return dpx, dpy, stops, rx, ry, (r_stop, ri, rj)


@pytest.mark.parametrize("nranks,midpt", [(1, True), (1, False), (2, True), (4, True)])
def test_departure_points_bitwise(dev, nranks, midpt):
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    dpx, dpy, stops, rx, ry, ref_stop = _departure_case(dev, grid, "full", 6, nranks, midpt)
    assert not ref_stop[0] and all(not s[0] for s in stops)
    assert np.array_equal(dpx, rx) and np.array_equal(dpy, ry)          # ghost ring included
    assert np.abs(rx).max() > 10.0 and rx[-1, 1:-1].any()                # metres; the fold filled the north ghost row


def test_departure_points_gx3_and_out_of_bounds(dev):
    """Displaced-pole grid (no fold); then a time step so long that departure points leave their cells: l_stop and the
    indices of the last offending cell as the reference reports them, first-order values on the physical cells."""
    grid = fx.make_grid("gx3")
    dpx, dpy, stops, rx, ry, ref_stop = _departure_case(dev, grid, "full", 6, 1, True)
    assert not ref_stop[0] and np.array_equal(dpx, rx) and np.array_equal(dpy, ry)
    dpx, dpy, stops, rx, ry, ref_stop = _departure_case(dev, grid, "full", 6, 1, True, dt_scale=4.0e3)
    assert ref_stop[0] and stops[0] == ref_stop and ref_stop[1] > 1 and ref_stop[2] > 1
    assert np.array_equal(phys(dpx), phys(rx)) and np.array_equal(phys(dpy), phys(ry))


def _once(dev):
    """The transport kernels are the same under both subcycle kernels: their host-build tests run with the product's."""
    if os.environ.get("EVP_B200_KERNEL") == "sweep":
        pytest.skip("transport kernels do not depend on the subcycle kernel; covered by the tma run")


# ---- horizontal transport, field reconstruction: make_masks + construct_fields -----------------
def _construct_case(dev, grid, scenario, nranks, geometry, seed=5):
    """construct_fields on the device against the oracle's (itself pinned to the reference text), every cell of every
    plane, ghost ring included."""
    case = fx.make_case(grid, scenario, ndte=6, dt=3600.0)
    d = fx.transport_tracers(grid, case.inp, seed, geometry)
    ncat = d["tm"].shape[0]
    if nranks == 1:
        e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
        try:
            e.set_transport_tracers(ncat, d["tracer_type"], d["depend"], d["has_dependents"], d)
            got = e.construct_fields(d["mm"], d["tm"])
            again = e.construct_fields(d["mm"], d["tm"])                  # no state carried from call to call
        finally:
            e.close()
        for k in got:
            assert np.array_equal(got[k], again[k]), k
    else:
        g = dev.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True)
        try:
            g.set_transport_tracers(ncat, d["tracer_type"], d["depend"], d["has_dependents"], d)
            got = g.construct_fields(d["mm"], d["tm"])
        finally:
            g.close()
    want = orc.construct_fields(case.dom, d["tracer_type"], d["depend"], d["has_dependents"], d["hm"], d["xav"], d["yav"],
                                d["xxav"], d["yyav"], d["mm"], d["tm"])
    return got, want, d


@pytest.mark.parametrize("nranks,geometry", [(1, "cice"), (1, "random"), (2, "random"), (4, "cice")])
def test_construct_fields_bitwise(dev, nranks, geometry):
    _once(dev)
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    got, want, d = _construct_case(dev, grid, "caps", nranks, geometry)
    for k in ("mc", "mx", "my", "tc", "tx", "ty"):
        assert got[k].shape == want[k].shape and np.array_equal(got[k], want[k]), k
        assert np.array_equal(np.signbit(got[k]), np.signbit(want[k])), k
    assert np.count_nonzero(want["tx"][:, 0]) > 200 and np.count_nonzero(want["tx"][:, 9]) > 20    # hice; hpnd on apnd
    assert want["tc"][:, :, -1, 1:-1].any()                              # the fold filled the north ghost row
    assert not np.array_equal(want["tc"][:, 0], d["tm"][:, 0])           # the centre value differs from the mean


def test_construct_fields_gx3_and_argument_errors(dev):
    _once(dev)
    grid = fx.make_grid("gx3")
    got, want, d = _construct_case(dev, grid, "full", 1, "random")
    for k in ("mc", "mx", "my", "tc", "tx", "ty"):
        assert np.array_equal(got[k], want[k]), k
    case = fx.make_case(grid, "full", ndte=6, dt=3600.0)
    e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
    try:
        with pytest.raises(dev.EvpError):                                # not set up yet
            e._rc_shape = (5, 11)
            e.construct_fields(d["mm"], d["tm"])
        bad = d["depend"].copy()
        bad[3] = 3                                                       # a type-2 tracer on Tsfc, which has no dependents
        with pytest.raises(dev.EvpError):
            e.set_transport_tracers(5, d["tracer_type"], bad, d["has_dependents"], d)
        bad = d["depend"].copy()
        bad[3] = 9                                                       # parent comes later in the table
        with pytest.raises(dev.EvpError):
            e.set_transport_tracers(5, d["tracer_type"], bad, d["has_dependents"], d)
        many = np.ones(dev.REMAP_MAX_TRACE + 1, dtype=np.int32)
        with pytest.raises(dev.EvpError):
            e.set_transport_tracers(5, many, 0 * many, 0 * many, d)
    finally:
        e.close()


# ---- all of horizontal_remap ---------------------------------------------------------------------
def _remap_case(dev, grid, scenario, nranks, dt_scale, geometry="cice", seed=7, ndte=6, tarear_scale=1.0, open_water=True):
    """evp(dt), then horizontal_remap on the device from the resident velocities against the oracle's (pinned to the
    reference text routine by routine and as a whole), every cell of mm and tm."""
    case = fx.make_case(grid, scenario, ndte=ndte, dt=3600.0)
    got = clone_case(case)
    d = fx.transport_tracers(grid, case.inp, seed, geometry)
    if not open_water:
        d["mm"][0] = 0.0                                    # no open water to run out of: an ice category does first
    ncat = d["tm"].shape[0]
    f = dict(grid.f)
    f["tarear"] = grid.f["tarear"] * tarear_scale           # > 1: exaggerated flux divergence, areas go negative
    dt = case.dt * dt_scale
    if nranks == 1:
        e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
        try:
            e.set_grid(f)
            e.set_transport_grid(f)
            e.set_transport_tracers(ncat, d["tracer_type"], d["depend"], d["has_dependents"], d)
            e.evp(case.dt, got.inp, got.io, got.out)
            mm, tm, stop = e.horizontal_remap(dt, d["mm"], d["tm"])
            stops = [stop]
        finally:
            e.close()
    else:
        g = dev.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True)
        try:
            g.set_grid(f)
            g.set_transport_grid(f)
            g.set_transport_tracers(ncat, d["tracer_type"], d["depend"], d["has_dependents"], d)
            g.evp(case.dt, got.inp, got.io, got.out)
            mm, tm, stops = g.horizontal_remap(dt, d["mm"], d["tm"])
        finally:
            g.close()
    geo = dict(d)
    geo.update({k: f[k] for k in ("dxu", "dyu", "HTN", "HTE", "tarear")})
    rmm, rtm = d["mm"].copy(), d["tm"].copy()
    rstop = orc.horizontal_remap(case.dom, dt, d["tracer_type"], d["depend"], d["has_dependents"], geo, got.io["uvel"],
                                 got.io["vvel"], rmm, rtm)
    return mm, tm, stops, rmm, rtm, rstop, d


@pytest.mark.parametrize("nranks,geometry,dt_scale", [(1, "cice", 40.0), (1, "random", 150.0), (2, "cice", 150.0), (4, "random", 40.0)])
def test_horizontal_remap_bitwise(dev, nranks, geometry, dt_scale):
    _once(dev)
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    mm, tm, stops, rmm, rtm, rstop, d = _remap_case(dev, grid, "caps", nranks, dt_scale, geometry)
    assert rstop == (0, 0, 0, 0) and all(s == (0, 0, 0, 0) for s in stops)
    assert np.array_equal(mm, rmm) and np.array_equal(tm, rtm)
    assert np.array_equal(np.signbit(tm), np.signbit(rtm))
    moved = np.abs(rmm - d["mm"])[:, 1:-1, 1:-1]
    assert moved.max() > 1e-3 and np.count_nonzero(moved) > 1000         # ice was transported
    assert np.abs(rtm - d["tm"])[:, 3:, 1:-1, 1:-1].max() > 1e-6           # the type-2 / type-3 tracers too
    tot0 = (d["mm"][:, 1:-1, 1:-1] * grid.f["tarea"][1:-1, 1:-1]).sum()    # area is conserved up to the clipping of
    tot1 = (rmm[:, 1:-1, 1:-1] * grid.f["tarea"][1:-1, 1:-1]).sum()        # negative areas and the open boundaries
    assert abs(tot1 - tot0) < 1e-3 * tot0


def test_horizontal_remap_gx3_and_stops(dev):
    """Displaced-pole grid; then a time step so long that the remap must stop: bad departure points, reported as the
    reference reports them."""
    _once(dev)
    grid = fx.make_grid("gx3")
    mm, tm, stops, rmm, rtm, rstop, d = _remap_case(dev, grid, "full", 1, 20.0, "random")
    assert rstop == (0, 0, 0, 0) and stops == [rstop]
    assert np.array_equal(mm, rmm) and np.array_equal(tm, rtm)
    grid = fx.synthetic_displaced(44, 26)                                 # the stop reports: a small displaced-pole grid
    mm, tm, stops, rmm, rtm, rstop, d = _remap_case(dev, grid, "full", 1, 4.0e3, "random")
    assert rstop[0] == 1 and stops == [rstop]
    # negative new areas: open water first, then (open water made thick enough) an ice category
    mm, tm, stops, rmm, rtm, rstop, d = _remap_case(dev, grid, "full", 1, 20.0, "random", tarear_scale=400.0)
    assert rstop[0] == 2 and rstop[2] > 1 and stops == [rstop]
    mm, tm, stops, rmm, rtm, rstop, d = _remap_case(dev, grid, "full", 1, 20.0, "random", tarear_scale=400.0, open_water=False)
    assert rstop[0] == 3 and rstop[1] >= 1 and rstop[2] > 1 and stops == [rstop]


# ---- transport_remap: the caller of horizontal_remap, from and to the ice state --------------------------------------
RHOS, LFRESH = 330.0, 2.835e6 - 2.501e6          # ice_constants.F90: rhos, Lfresh = Lsub - Lvap


def _transport_state(grid, case, seed):
    """aice0, aicen, vicen, vsnon, trcrn consistent with fx.transport_tracers (tracers 3.. of the table are trcrn; the
    snow enthalpy tracer qsno1 = trcrn 6 is stored as the model stores it, without the rhos * Lfresh shift)."""
    d = fx.transport_tracers(grid, case.inp, seed, "cice")
    mm, tm = d["mm"], d["tm"]
    st = {"aice0": mm[0].copy(), "aicen": mm[1:].copy(), "vicen": mm[1:] * tm[:, 0], "vsnon": mm[1:] * tm[:, 1],
          "trcrn": np.ascontiguousarray(tm[:, 2:])}
    return d, st


def _transport_case(dev, grid, scenario, nranks, dt_scale, seed=9):
    case = fx.make_case(grid, scenario, ndte=6, dt=3600.0)
    got = clone_case(case)
    d, st = _transport_state(grid, case, seed)
    ncat = st["aicen"].shape[0]
    f = grid.f
    dt = case.dt * dt_scale
    par = (6, 1, RHOS, LFRESH)                                            # nt_qsno, nslyr, rhos, Lfresh
    if nranks == 1:
        e = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True)
        try:
            e.set_grid(f)
            e.set_transport_grid(f)
            e.set_transport_tracers(ncat, d["tracer_type"], d["depend"], d["has_dependents"], d)
            e.evp(case.dt, got.inp, got.io, got.out)
            new, stop = e.transport_remap(dt, st, *par)
            stops = [stop]
        finally:
            e.close()
    else:
        g = dev.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True)
        try:
            g.set_grid(f)
            g.set_transport_grid(f)
            g.set_transport_tracers(ncat, d["tracer_type"], d["depend"], d["has_dependents"], d)
            g.evp(case.dt, got.inp, got.io, got.out)
            new, stops = g.transport_remap(dt, st, *par)
        finally:
            g.close()
    geo = dict(d)
    geo.update({k: f[k] for k in ("dxu", "dyu", "HTN", "HTE", "tarear")})
    ref = {k: v.copy() for k, v in st.items()}
    rstop = orc.transport_remap(case.dom, dt, d["tracer_type"], d["depend"], d["has_dependents"], geo, got.io["uvel"],
                                got.io["vvel"], ref, *par)
    return new, stops, ref, rstop, st


@pytest.mark.parametrize("nranks", [1, 2, 4])
def test_transport_remap_bitwise(dev, nranks):
    _once(dev)
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    new, stops, ref, rstop, st = _transport_case(dev, grid, "caps", nranks, 100.0)
    assert rstop == (0, 0, 0, 0) and all(s == rstop for s in stops)
    for k in ("aice0", "aicen", "vicen", "vsnon", "trcrn"):
        assert np.array_equal(new[k], ref[k]), k                          # ghost cells too (bound_state), except aice0's
        assert np.array_equal(np.signbit(new[k]), np.signbit(ref[k])), k
    assert np.abs(ref["aicen"] - st["aicen"])[:, 1:-1, 1:-1].max() > 1e-3 and np.abs(ref["vicen"] - st["vicen"]).max() > 1e-3
    assert ref["aicen"][:, -1, 1:-1].any()                                # the fold filled the north ghost row


# ---- dynamics + advection as one resident phase: evp and transport_remap step after step -------------------------------
def _dynamics_steps(dev, grid, nsteps, nranks, scenario="caps", ndte=4, dt_scale=20.0, seed=11):
    """step_dynamics' two device calls, `nsteps` times (source/ice_step_mod.F90:1113-1130): evp from the aggregated ice
    state, transport_remap of the state from the velocities evp left on the device.  The oracle does the same with its
    own state; after every step the dynamic state of evp and the whole ice state are compared bit for bit."""
    case = fx.make_case(grid, scenario, ndte=ndte, dt=3600.0)
    d, st0 = _transport_state(grid, case, seed)
    ncat = st0["aicen"].shape[0]
    f = grid.f
    par = (6, 1, RHOS, LFRESH)
    geo = dict(d)
    geo.update({k: f[k] for k in ("dxu", "dyu", "HTN", "HTE", "tarear")})

    def evp_inputs(st, step):
        s = {"aice": st["aicen"].sum(axis=0), "vice": st["vicen"].sum(axis=0), "vsno": st["vsnon"].sum(axis=0)}
        s["aice_init"] = s["aice"].copy()
        s["strength"] = fx.PSTAR * s["vice"] * np.exp(-fx.CSTAR * (1.0 - s["aice"]))     # Hibler (ice_mechred.F90:2263)
        inp = dict(s)
        inp.update(fx.forcing(grid, s, step, steps_per_cycle=24))
        return inp

    dev_st = {k: v.copy() for k, v in st0.items()}
    ref_st = {k: v.copy() for k, v in st0.items()}
    got, ref = clone_case(case), clone_case(case)
    single = nranks == 1
    h = dev.EvpGpu.for_domain(case.dom, case.ndte, case.dt, strict=True) if single else \
        dev.EvpGroup(case.dom, case.ndte, case.dt, nranks, strict=True)
    moved = 0.0
    try:
        h.set_grid(f)
        h.set_transport_grid(f)
        h.set_transport_tracers(ncat, d["tracer_type"], d["depend"], d["has_dependents"], d)
        for step in range(nsteps):
            got.inp, ref.inp = evp_inputs(dev_st, step), evp_inputs(ref_st, step)
            got.out, ref.out = fx.zeroed_outputs(case.dom), fx.zeroed_outputs(case.dom)
            h.evp(case.dt, got.inp, got.io, got.out)
            orc.run_case(ref)
            new, stop = h.transport_remap(case.dt * dt_scale, dev_st, *par)
            rstop = orc.transport_remap(case.dom, case.dt * dt_scale, d["tracer_type"], d["depend"], d["has_dependents"], geo,
                                        ref.io["uvel"], ref.io["vvel"], ref_st, *par)
            stops = [stop] if single else stop
            assert rstop == (0, 0, 0, 0) and all(s == rstop for s in stops), (step, stops, rstop)
            for k in ("uvel", "vvel") + fx.STRESS_FIELDS:
                assert np.array_equal(phys(got.io[k]), phys(ref.io[k])), (step, k)
            for k in ref_st:
                assert np.array_equal(new[k], ref_st[k]), (step, k)
            moved = max(moved, float(np.abs(new["aicen"] - dev_st["aicen"])[:, 1:-1, 1:-1].max()))
            dev_st = new
    finally:
        h.close()
    return moved, float(np.abs(ref_st["aicen"] - st0["aicen"])[:, 1:-1, 1:-1].max())


@pytest.mark.parametrize("nranks", [1, 2])
def test_dynamics_and_advection_steps_bitwise(dev, nranks):
    _once(dev)
    grid = fx.synthetic_tripole(96, 40, wet_poles=True)
    moved, total = _dynamics_steps(dev, grid, 6, nranks)
    assert moved > 1e-4 and total > 1e-3
