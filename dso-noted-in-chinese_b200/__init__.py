"""dso_b200 -- B200 (sm_100a) implementation of DSO's photometric Gauss-Newton hot path.

The product is the shared library `libdso_b200.so` (CUDA kernels behind the C ABI of include/dso_b200.h)
plus the C++ classes in host/ that keep the reference's member-function signatures.  The Python modules
here are plumbing: `capi` binds the C ABI through ctypes for the tests and bench.py, `synthetic` generates
the seeded workloads.  The directory name contains dashes, so load it with
`__graft_entry__.load_package()` (registers the package as `dso_b200`).
"""
from . import capi, synthetic  # noqa: F401
from .capi import Context, DsoError  # noqa: F401
