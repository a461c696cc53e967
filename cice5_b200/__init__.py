"""cice5_b200 -- B200-native EVP momentum solver, a drop-in for CICE5's ``evp(dt)``.

Only what the hot path needs lives here: ``csrc/`` (sm_100a CUDA kernels + the C ABI of
``include/evp_b200.h``), ``evp.py`` (host-side mirror of the reference interface over ctypes),
``fixtures.py`` (synthetic grids / ice state / forcing) and ``build.py`` (nvcc driver).
"""
__version__ = "0.1.0"
