"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/cpprl_b200.h declares,
and refuses (loudly, no fallback) to compute without an sm_100 device."""
import ctypes as C
import re
from pathlib import Path

import pytest
import torch

from pytorch_cpp_rl_b200 import cabi

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "cpprl_b200.h"


def declared_functions():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(cpprl_b200_\w+)\s*\(", text)))


def test_header_declares_the_path():
    names = declared_functions()
    for must in ("cpprl_b200_compute_returns", "cpprl_b200_advantages", "cpprl_b200_gather_rows", "cpprl_b200_rms_update",
                 "cpprl_b200_normalize_observations", "cpprl_b200_adam_step", "cpprl_b200_rmsprop_step",
                 "cpprl_b200_evaluate_actions", "cpprl_b200_act", "cpprl_b200_ppo_update", "cpprl_b200_a2c_update", "cpprl_b200_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(str(cabi.LIB_PATH))
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, f"declared in include/cpprl_b200.h but not exported: {missing}"


def test_abi_version_and_struct_sizes():
    lib = cabi.load()
    assert lib.cpprl_b200_abi_version() == 1
    # the ctypes mirrors must have the C layout (LP64): see include/cpprl_b200.h
    assert C.sizeof(cabi.Policy) == 8 + 8 + 24 + 8 + 8 + 8 + 8 + 4 * 8
    assert C.sizeof(cabi.Rollout) == 4 * 8 + 7 * 8
    assert C.sizeof(cabi.PPOOptions) == 9 * 4
    assert C.sizeof(cabi.PPOStats) == 7 * 4


@pytest.mark.skipif(torch.cuda.is_available(), reason="needs a box WITHOUT a GPU")
def test_no_cpu_fallback():
    lib = cabi.load()
    buf = (C.c_float * 16)()
    rc = lib.cpprl_b200_compute_returns(buf, buf, buf, buf, buf, 1, 1, 1, C.c_float(0.99), C.c_float(0.95), None)
    assert rc == cabi.ECUDA
    assert b"no CPU fallback" in lib.cpprl_b200_last_error() or b"CUDA" in lib.cpprl_b200_last_error()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        from pytorch_cpp_rl_b200 import api
        api.gather_rows(torch.zeros(4, 4), torch.zeros(2, dtype=torch.int64))
