"""Horizontal transport, first stages on the device, bit for bit against the oracle
(oracle/remap_oracle.c, itself pinned to the reference text): departure_points (source/ice_transport_remap.F90:1493-1670)
+ the ghost update of dpx, dpy (:574-578) from the resident velocities of the last evp; make_masks + construct_fields +
limited_gradient (:867-1484) + the ghost updates of mc, mx, my, tc, tx, ty (:579-584, :616-621); and all of horizontal_remap (:309-850: the two stages
above, then locate_triangles, triangle_coordinates, transport_integrals per edge and update_fields).  The test bodies are the
ones the CPU tier runs on the host build of the device sources."""
import pytest

from cice5_b200 import evp
from tests import test_emu_device_sources as shared

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.mark.parametrize("nranks,midpt", [(1, True), (1, False), (2, True), (4, True)])
def test_departure_points_bitwise_on_device(nranks, midpt):
    shared.test_departure_points_bitwise(evp, nranks, midpt)


def test_departure_points_gx3_and_out_of_bounds_on_device():
    shared.test_departure_points_gx3_and_out_of_bounds(evp)


def test_departure_points_one_degree_tripole_on_device():
    from cice5_b200 import fixtures as fx
    import numpy as np
    grid = fx.make_grid("tp1")
    dpx, dpy, stops, rx, ry, ref_stop = shared._departure_case(evp, grid, "caps", 120, 1, True)
    assert not ref_stop[0] and np.array_equal(dpx, rx) and np.array_equal(dpy, ry)


@pytest.mark.parametrize("nranks,geometry", [(1, "cice"), (1, "random"), (2, "random"), (4, "cice")])
def test_construct_fields_bitwise_on_device(nranks, geometry):
    shared.test_construct_fields_bitwise(evp, nranks, geometry)


def test_construct_fields_gx3_and_argument_errors_on_device():
    shared.test_construct_fields_gx3_and_argument_errors(evp)


def test_construct_fields_one_degree_tripole_on_device():
    """The 1-degree tripole with the full tracer table: 6 + 55 input planes, 18 + 165 output planes."""
    from cice5_b200 import fixtures as fx
    import numpy as np
    grid = fx.make_grid("tp1")
    got, want, d = shared._construct_case(evp, grid, "caps", 1, "cice")
    for k in ("mc", "mx", "my", "tc", "tx", "ty"):
        assert np.array_equal(got[k], want[k]), k
        assert np.array_equal(np.signbit(got[k]), np.signbit(want[k])), k
    assert np.count_nonzero(want["tx"]) > 10000


@pytest.mark.parametrize("nranks,geometry,dt_scale", [(1, "cice", 40.0), (1, "random", 150.0), (2, "cice", 150.0), (4, "random", 40.0)])
def test_horizontal_remap_bitwise_on_device(nranks, geometry, dt_scale):
    shared.test_horizontal_remap_bitwise(evp, nranks, geometry, dt_scale)


def test_horizontal_remap_gx3_and_stops_on_device():
    shared.test_horizontal_remap_gx3_and_stops(evp)


def test_horizontal_remap_one_degree_tripole_on_device():
    """The 1-degree tripole after a full evp (ndte = 120), a transport step of six hours."""
    from cice5_b200 import fixtures as fx
    import numpy as np
    grid = fx.make_grid("tp1")
    mm, tm, stops, rmm, rtm, rstop, d = shared._remap_case(evp, grid, "caps", 1, 6.0, "cice", ndte=120)
    assert rstop == (0, 0, 0, 0) and stops == [rstop]
    assert np.array_equal(mm, rmm) and np.array_equal(tm, rtm) and np.array_equal(np.signbit(tm), np.signbit(rtm))
    assert np.abs(rmm - d["mm"]).max() > 1e-4


@pytest.mark.parametrize("nranks", [1, 2, 4])
def test_transport_remap_bitwise_on_device(nranks):
    """transport_remap (source/ice_transport_driver.F90:198-627): state_to_tracers, horizontal_remap, tracers_to_state,
    bound_state, from and to aice0 / aicen / vicen / vsnon / trcrn."""
    shared.test_transport_remap_bitwise(evp, nranks)


@pytest.mark.parametrize("nranks", [1, 2])
def test_dynamics_and_advection_steps_bitwise_on_device(nranks):
    """evp + transport_remap, six steps in a row on the device-resident state."""
    shared.test_dynamics_and_advection_steps_bitwise(evp, nranks)


def test_dynamics_and_advection_one_degree_tripole_on_device():
    """The 1-degree tripole: 12 steps of evp (ndte = 120) + transport_remap, every step bit for bit against the oracle."""
    from cice5_b200 import fixtures as fx
    grid = fx.make_grid("tp1")
    moved, total = shared._dynamics_steps(evp, grid, 12, 1, scenario="caps", ndte=120, dt_scale=2.0)
    assert total > 1e-4
