"""Shared by the EAP tests (host build and device): an EAP state, and the comparison of device and oracle."""
import numpy as np

from cice5_b200 import fixtures as fx
from oracle import oracle as orc
from tests.test_gpu_parity import OUT_CHECK
from tests.util import clone_case, compare_cases

EAP_FIELDS = orc.EAP_CORNER + orc.EAP_AVG + orc.EAP_OUT


def eap_state(dom, seed=5, fresh=False):
    """fresh: as init_eap leaves it (isotropic, a11_k = 0.5); else a seeded valid structure tensor (eigenvalues in
    (0, 1): a12^2 < a11 a22) on every cell, ghost ring included, and garbage in the diagnostics (the call zeroes them)."""
    e = orc.eap_state(dom)
    if fresh:
        return e
    r = np.random.default_rng(seed)
    sh = e["a11"].shape
    for c in (1, 2, 3, 4):
        e[f"a11_{c}"] = r.uniform(0.3, 0.9, sh)
        e[f"a12_{c}"] = r.uniform(-0.9, 0.9, sh) * np.sqrt(e[f"a11_{c}"] * (1.0 - e[f"a11_{c}"]))
    e["a11"], e["a12"] = r.uniform(0.3, 0.9, sh), r.uniform(-0.2, 0.2, sh)
    return e


def clone_eap(e):
    return {k: v.copy() for k, v in e.items()}


def compare_eap(got_case, ref_case, got_eap, ref_eap, dom):
    """Every inout / out field of the call and every EAP field on the cells both sides define, bit for bit."""
    compare_cases(got_case, ref_case, fields_out=OUT_CHECK, bitwise=True)
    bad = []
    for n in EAP_FIELDS:
        a, b = got_eap[n], ref_eap[n]
        if n in orc.EAP_CORNER:              # persistent state incl. the north / east ghost cells of the T list
            a, b = a[1:, 1:], b[1:, 1:]
        else:
            a, b = a[1:dom.ny + 1, 1:dom.nx + 1], b[1:dom.ny + 1, 1:dom.nx + 1]
        if not np.array_equal(a, b):
            bad.append((n, int(np.count_nonzero(a != b)), float(np.abs(a - b).max())))
    assert not bad, f"EAP fields differing from the oracle (name, cells, max abs): {bad}"
