"""Horizontal transport, first stage on the device: departure_points
(source/ice_transport_remap.F90:1493-1670) + the ghost update of dpx, dpy (:574-578) from the resident velocities of the
last evp, bit for bit against the oracle (oracle/remap_oracle.c, itself pinned to the reference text).  The test bodies
are the ones the CPU tier runs on the host build of the device sources."""
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
