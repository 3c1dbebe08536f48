"""Kernel-level check of the tcgen05 3xTF32 GEMM (csrc/cnn_gemm_tc.cuh) through the C-ABI test hook
cpprl_b200_debug_tc_gemm: C = A B with plain row-major operands against a float64 product.

The bar is fp32-grade accuracy at EVERY reduction length (the chunked accumulation exists because the tensor core's
truncating fp32 adds cost 1e-8 * K otherwise): relative error <= 1.5e-6 of max|C|, i.e. no worse than an fp32 FFMA GEMM."""
import ctypes as C

import pytest
import torch

from pytorch_cpp_rl_b200 import cabi

pytestmark = pytest.mark.gpu

CASES = [(bn, a_t, 256, bn, 96) for bn in (32, 64, 128) for a_t in (0, 1)] + [
    (32, 0, 200, 32, 160),        # ragged M
    (64, 1, 576, 64, 1000),       # K not a multiple of the 32-wide k tile, A given transposed (weight-gradient form)
    (128, 0, 1000, 512, 1568),    # the linear layer's forward shape
    (128, 1, 1568, 512, 4096),    # long reduction: the truncation error would be 4e-5 without chunked accumulation
    (32, 0, 40000, 32, 256),      # conv1-like: many row tiles, 32 output channels
    (64, 0, 20736, 64, 512),      # conv2-like
]


@pytest.mark.parametrize("bn,a_t,M,N,K", CASES)
def test_tc_gemm_matches_float64(bn, a_t, M, N, K):
    lib = cabi.load()
    g = torch.Generator(device="cuda").manual_seed(1000 * bn + M + K)
    A = torch.randn(M, K, device="cuda", generator=g)
    B = torch.randn(K, N, device="cuda", generator=g)
    a_in = A.t().contiguous() if a_t else A
    out = torch.full((M, N), float("nan"), device="cuda")
    rc = lib.cpprl_b200_debug_tc_gemm((bn << 4) | a_t, C.c_void_p(a_in.data_ptr()), C.c_void_p(B.data_ptr()),
                                      C.c_void_p(out.data_ptr()), M, N, K, None)
    assert rc == 0, cabi.last_error() if hasattr(cabi, "last_error") else rc
    torch.cuda.synchronize()
    ref = A.double() @ B.double()
    assert not torch.isnan(out).any()
    err = ((out.double() - ref).abs().max() / ref.abs().max()).item()
    assert err <= 1.5e-6, err
