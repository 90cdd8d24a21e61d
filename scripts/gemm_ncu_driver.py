"""Two 4096^3 GEMMs through the pipelined DMMA kernel: (A m-fast, B k-fast) then (A k-fast, B k-fast); ncu target."""
import sys, torch
sys.path.insert(0, '.')
import tci_b200 as T
from tci_b200.dense import Mat, gemm
h = T.default_handle(); dev = "cuda:0"
M = N = K = 4096
B = torch.randn(N, K, device=dev, dtype=torch.float64); Bm = Mat(B, K, N)
Cm = Mat.empty(M, N, dev)
for ta in (False, True, False, True):
    A = torch.randn((M, K) if ta else (K, M), device=dev, dtype=torch.float64)
    Am = Mat(A, K, M) if ta else Mat(A, M, K)
    gemm(h, Am, Bm, Cm, transa=ta)
torch.cuda.synchronize()
print("ok")
