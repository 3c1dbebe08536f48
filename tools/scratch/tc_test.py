import sys, ctypes as C
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import torch
from pytorch_cpp_rl_b200 import cabi
lib = cabi.load()
torch.manual_seed(0)
def run(variant, M, N, K, time_it=False):
    a_t, bn = variant & 1, variant >> 4
    A = torch.randn(M, K, device="cuda"); B = torch.randn(K, N, device="cuda")
    Ain = A.t().contiguous() if a_t else A
    Cout = torch.full((M, N), float("nan"), device="cuda")
    args = (variant, C.c_void_p(Ain.data_ptr()), C.c_void_p(B.data_ptr()), C.c_void_p(Cout.data_ptr()), M, N, K, None)
    rc = lib.cpprl_b200_debug_tc_gemm(*args)
    torch.cuda.synchronize()
    ref = (A.double() @ B.double())
    err = ((Cout.double() - ref).abs().max() / ref.abs().max()).item()
    torch.backends.cuda.matmul.allow_tf32 = False
    ref32 = ((A @ B).double() - ref).abs().max().item() / ref.abs().max().item()
    msg = ""
    if time_it:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            lib.cpprl_b200_debug_tc_gemm(*args)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        msg = f" {ms*1e3:.0f} us = {2*M*N*K/ms/1e9:.1f} TFLOP/s"
    print(f"a_t={a_t} bn={bn} M={M} N={N} K={K}: rc={rc} rel err {err:.2e} (torch fp32 {ref32:.2e}) nan={torch.isnan(Cout).sum().item()}{msg}", flush=True)
for bn in (32, 64, 128):
    for a_t in (0, 1):
        run((bn << 4) | a_t, 256, bn, 96)
run((32 << 4) | 0, 200, 32, 160); run((64 << 4) | 1, 576, 64, 1000)
run((128 << 4) | 0, 1000, 512, 1568, True); run((128 << 4) | 1, 1568, 512, 4096, True)
run((32 << 4) | 0, 1638400, 32, 256, True); run((64 << 4) | 0, 331776, 64, 512, True)
