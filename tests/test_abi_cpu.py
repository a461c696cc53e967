"""The C-ABI library loads and exports every symbol include/evp_b200.h declares; without a GPU
it must fail loudly (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from cice5_b200 import build, evp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    txt = open(os.path.join(ROOT, "include", "evp_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(evp_gpu_\w+)\s*\(", txt)))


@pytest.mark.parametrize("strict", [False, True])
def test_library_exports_every_declared_symbol(strict):
    names = _declared_functions()
    assert len(names) >= 15
    lib = evp.load_library(strict)
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(evp.EXPORTS)


def test_strerror_and_status_codes():
    lib = evp.load_library()
    assert lib.evp_gpu_strerror(0) == b"ok"
    for rc in range(1, 7):
        assert lib.evp_gpu_strerror(rc).startswith(b"evp_gpu:")


def test_config_struct_matches_header():
    txt = open(os.path.join(ROOT, "include", "evp_b200.h")).read()
    body = re.search(r"typedef struct \{(.*?)\} evp_config_t;", txt, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        typ, rest = decl.split(None, 1)
        fields += [(n.strip(), typ) for n in rest.split(",")]

    def cname(n, t):
        if t is C.c_int:
            return (n, "int")
        if t is C.c_double:
            return (n, "double")
        return (f"{n}[{t._length_}]", "int")            # int array member
    got = [cname(n, t) for n, t in evp.Config._fields_]
    assert got == fields


def test_no_gpu_means_loud_failure(have_gpu):
    if have_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(evp.EvpError, match="no usable CUDA device|no CPU fallback"):
        evp.EvpGpu(100, 116, 120, 3600.0)


def test_argument_validation_precedes_device_probe():
    with pytest.raises(evp.EvpError, match="unsupported"):
        evp.EvpGpu(100, 116, 120, 3600.0, kdyn=3)
    with pytest.raises(evp.EvpError, match="unsupported"):
        evp.EvpGpu(100, 116, 120, 3600.0, kdyn=2, revised_evp=True)      # not a configuration of the reference's eap
    with pytest.raises(evp.EvpError, match="invalid argument"):
        # revised EVP split over tasks needs the global xmin (ice_dyn_shared.F90:220-223)
        evp.EvpGpu(100, 116, 120, 3600.0, revised_evp=True, nranks=2, rank=0)
    with pytest.raises(evp.EvpError):
        evp.EvpGpu(100, 116, 120, 3600.0, nranks=3)      # 3 does not divide 100


def test_product_path_never_imports_oracle():
    import sys
    for mod in ("cice5_b200.evp", "cice5_b200.fixtures", "cice5_b200.build"):
        src = open(sys.modules[mod].__file__ if mod in sys.modules else
                   os.path.join(ROOT, *mod.split(".")) + ".py").read()
        assert "import oracle" not in src and "from oracle" not in src


def test_header_is_plain_c():
    """The boundary is a C ABI (bound from Fortran iso_c_binding): the header must compile as C99."""
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    r = subprocess.run([cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c",
                        os.path.join(ROOT, "include", "evp_b200.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_slab_start_validation_precedes_device_probe(have_gpu):
    """Unequal x-slabs (evp_config_t.slab_start): inconsistent boundaries are argument errors, a
    consistent set gets as far as the device probe."""
    with pytest.raises(evp.EvpError, match="first column per rank"):
        evp.EvpGpu(100, 116, 120, 3600.0, nranks=3, rank=0, slab_start=[0, 40])            # wrong length
    with pytest.raises(evp.EvpError, match="unsupported"):
        evp.EvpGpu(100, 116, 120, 3600.0, nranks=3, rank=1, slab_start=[0, 40, 41])        # 1-column slab
    if not have_gpu:
        with pytest.raises(evp.EvpError, match="no usable CUDA device|no CPU fallback"):
            evp.EvpGpu(100, 116, 120, 3600.0, nranks=3, rank=2, slab_start=[0, 30, 71])    # 3 does not divide 100
