"""Shared helpers for the parity tests: error norms in the form BASELINE.json asks for
(relative to the field's own magnitude) and case cloning."""
import copy

import numpy as np

from cice5_b200 import fixtures as fx

STATE = ("uvel", "vvel") + fx.STRESS_FIELDS
REL_TOL = 1.0e-10   # BASELINE.json north_star: "within 1e-10 relative (FP64)"


def clone_case(case):
    c = copy.copy(case)
    c.inp = {k: v.copy() for k, v in case.inp.items()}
    c.io = {k: v.copy() for k, v in case.io.items()}
    c.out = {k: v.copy() for k, v in case.out.items()}
    return c


def phys(a):
    return a[1:-1, 1:-1]


def rel_errors(x, ref):
    """(err_inf, err_rms) = (max|x-ref| / max|ref|, rms(x-ref) / rms(ref)) on physical cells."""
    x, ref = phys(x).astype("f8"), phys(ref).astype("f8")
    d = x - ref
    den_inf = np.abs(ref).max()
    den_rms = np.sqrt((ref * ref).mean())
    if den_inf == 0.0:
        return (0.0, 0.0) if np.abs(d).max() == 0.0 else (np.inf, np.inf)
    return np.abs(d).max() / den_inf, np.sqrt((d * d).mean()) / den_rms


def compare_cases(got, ref, fields_io=None, fields_out=None, tol=REL_TOL, bitwise=False,
                  ghosts=False):
    """Assert that `got` matches `ref` on every inout / out field; returns {field: (inf, rms)}."""
    rep = {}
    bad = []
    for group, names in (("io", fields_io), ("out", fields_out)):
        a, b = getattr(got, group), getattr(ref, group)
        for n in (names if names is not None else a.keys()):
            if b.get(n) is None or a.get(n) is None:
                continue
            x, r = a[n], b[n]
            if bitwise:
                xs, rs = (x, r) if ghosts else (phys(x), phys(r))
                same = np.array_equal(xs, rs)
                rep[n] = (0.0, 0.0) if same else rel_errors(x, r)
                if not same:
                    bad.append((n, rep[n], int((xs != rs).sum())))
            else:
                rep[n] = rel_errors(x, r)
                if not (rep[n][0] <= tol):
                    bad.append((n, rep[n]))
    assert not bad, f"parity failure ({'bitwise' if bitwise else tol}): {bad}"
    return rep
