import sys, ctypes as C
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import torch
from pytorch_cpp_rl_b200 import cabi
lib = cabi.load()
torch.set_printoptions(linewidth=200, sci_mode=False)
M, N, K = 128, 32, 8
def go(variant, A, B):
    a_mn, b_mn = variant & 1, variant & 2
    Ain = A.t().contiguous() if a_mn else A.contiguous()
    Bin = B.contiguous() if b_mn else B.t().contiguous()
    Cout = torch.full((M, N), float("nan"), device="cuda")
    lib.cpprl_b200_debug_tc_gemm(variant, C.c_void_p(Ain.data_ptr()), C.c_void_p(Bin.data_ptr()), C.c_void_p(Cout.data_ptr()), M, N, K, None)
    torch.cuda.synchronize()
    return Cout
# probe A (MN-major) with B K-major selector: C[m][n] = A_hw(m, n % 8)
A = (torch.arange(M, device="cuda").float()[:, None] * 8 + torch.arange(K, device="cuda").float()[None, :])
Bsel = torch.zeros(K, N, device="cuda")
for n in range(N):
    Bsel[n % 8, n] = 1
for mode in (0, 4):
    Cout = go((32 << 4) | 1 | mode, A, Bsel)
    print("A MN-major, mode", mode >> 2, "rows 0..11 cols 0..7 (expect m*8+k):")
    print(Cout[:12, :8].int())
    print("rows 32..35:", Cout[32:36, :8].int().tolist())
# probe B (MN-major) with A K-major selector: A[m][k] = 1 if k == m % 8 -> C[m][n] = B_hw(m % 8, n)
Bv = (torch.arange(K, device="cuda").float()[:, None] * 32 + torch.arange(N, device="cuda").float()[None, :])
Asel = torch.zeros(M, K, device="cuda")
for m in range(M):
    Asel[m, m % 8] = 1
for mode in (0, 4):
    Cout = go((32 << 4) | 2 | mode, Asel, Bv)
    print("B MN-major, mode", mode >> 2, "rows 0..7 (expect k*32+n):")
    print(Cout[:8, :16].int())
