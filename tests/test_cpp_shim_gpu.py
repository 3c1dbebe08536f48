"""Runs the C++ drop-in test program (tests/cpp/host_shim_test.cpp): the reference's class API over the C ABI,
driven like the reference's own doctest cases and checked against the compiled reference."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
EXE = ROOT / "tests" / "cpp" / "host_shim_test"


@pytest.mark.gpu
def test_cpp_drop_in_against_reference():
    if not EXE.exists():
        pytest.skip("tests/cpp/host_shim_test not built (run tests/cpp/build.sh where oracle/_ref exists)")
    res = subprocess.run([str(EXE)], capture_output=True, text=True, timeout=600)
    print(res.stdout[-6000:])
    print(res.stderr[-2000:])
    assert res.returncode == 0, "C++ drop-in checks failed"


def test_host_library_links():
    """CPU: the drop-in shared library exists, loads, and defines the reference's class methods."""
    lib = ROOT / "pytorch_cpp_rl_b200" / "lib" / "libcpprl_host.so"
    if not lib.exists():
        pytest.skip("libcpprl_host.so not built")
    import ctypes
    import torch  # noqa: F401  (libtorch must be resident first)
    ctypes.CDLL(str(ROOT / "pytorch_cpp_rl_b200" / "lib" / "libcpprl_b200.so"), mode=ctypes.RTLD_GLOBAL)
    ctypes.CDLL(str(lib))
    syms = subprocess.run(["nm", "-DC", "--defined-only", str(lib)], capture_output=True, text=True).stdout
    for must in ("cpprl::PPO::update(cpprl::RolloutStorage&, float)", "cpprl::A2C::update(cpprl::RolloutStorage&, float)",
                 "cpprl::RolloutStorage::compute_returns(at::Tensor, bool, float, float)",
                 "cpprl::PolicyImpl::evaluate_actions(at::Tensor, at::Tensor, at::Tensor, at::Tensor) const",
                 "cpprl::FeedForwardGenerator::next()", "cpprl::RecurrentGenerator::next()",
                 "cpprl::RunningMeanStdImpl::update(at::Tensor)",
                 "cpprl::ObservationNormalizerImpl::process_observation(at::Tensor) const"):
        assert must in syms, must
