"""cpprl-b200: B200-native on-policy update hot path of CppRl (pytorch-cpp-rl).

Only what the path needs lives here: `csrc/` (sm_100a kernels + the C ABI declared in
include/cpprl_b200.h), `host/` (the C++ drop-in shim keeping the reference's `cpprl::` classes),
and a thin Python mirror (`api`) used by the tests and bench.  Importing the package does not load
the native library; the first call does, and it fails loudly if the library has not been built.
"""
__all__ = ["api", "cabi", "workloads", "build"]
