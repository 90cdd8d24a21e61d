"""Small driver for ncu: three 5120 x 5120 Pi fills (Lorentzian functor) and three 4096^3 DMMA GEMMs."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import tci_b200 as T
from tci_b200.dense import Mat, gemm
h = T.default_handle(); dev = "cuda:0"
ld = [20] * 20
f = T.LorentzFunction(ld)
rng = np.random.default_rng(0)
left = rng.integers(1, 21, size=(5120, 10)).astype(np.int32); right = rng.integers(1, 21, size=(5120, 10)).astype(np.int32)
buf = torch.empty(5120 * 5120, dtype=torch.float64, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for it in range(4):
    torch.cuda.synchronize(); e0.record()
    f.evaluate_batch(left, right, 0, out=buf, stride_c=5120, stride_j=5120)
    e1.record(); torch.cuda.synchronize()
    if it: print(f"K1 fill 5120^2: {e0.elapsed_time(e1)*1e3:.1f} us -> {8*5120*5120/e0.elapsed_time(e1)/1e6:.0f} GB/s written, {5120*5120/e0.elapsed_time(e1)/1e6:.1f} Gevals/s")
A = Mat(torch.randn(4096, 4096, device=dev, dtype=torch.float64), 4096, 4096); B = Mat(torch.randn(4096, 4096, device=dev, dtype=torch.float64), 4096, 4096)
C = Mat.empty(4096, 4096, dev)
for it in range(4):
    torch.cuda.synchronize(); e0.record(); gemm(h, A, B, C); e1.record(); torch.cuda.synchronize()
    if it: print(f"GEMM 4096^3: {e0.elapsed_time(e1):.3f} ms -> {2*4096**3/e0.elapsed_time(e1)/1e9:.2f} TFLOP/s")
