"""Resident-state protocol of evp_gpu_run (evp_gpu_set_resident / invalidate_state) and the restart hand-off
(evp_gpu_restart_fixup: source/ice_restart_driver.F90:370-394, 431-466) on the device: the same test bodies that
run on the host build of the device sources (tests/test_emu_device_sources.py), here against the CUDA library."""
import pytest

from cice5_b200 import evp
from tests import test_emu_device_sources as shared

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.mark.parametrize("nranks", [1, 2, 4])
def test_restart_fixup_bitwise_on_device(nranks):
    shared.test_restart_fixup_bitwise(evp, nranks)


def test_resident_protocol_on_device():
    shared.test_resident_protocol_of_evp_gpu_run(evp)
