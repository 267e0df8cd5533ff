"""vpt-b200: B200-native Yocto/Pathtrace sample loop (pathtrace_samples -> shade_*).

The package directory name contains a hyphen, so it is registered under the
importable name ``vpt_b200`` by ``__graft_entry__.load_package()``.
Product code is C++/CUDA (csrc/, host/); the Python here is the ctypes binding
and the fixture tooling used by tests/ and bench.py.
"""
from . import archive, binding, sharding  # noqa: F401
