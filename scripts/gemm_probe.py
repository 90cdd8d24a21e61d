"""tci_gemm_f64 (hand-written DMMA kernel) vs cuBLAS DGEMM (torch.matmul) as the measured FP64 tensor peak; run under gpurun."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import tci_b200 as T
from tci_b200.dense import Mat, gemm
h = T.default_handle()
dev = "cuda:0"
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3
for (M, N, K) in [(4096, 4096, 4096), (8192, 8192, 2048), (40000, 800, 200), (3200, 3200, 3200), (256, 5120, 256), (200, 800, 200)]:
    A = torch.randn(K, M, device=dev, dtype=torch.float64); B = torch.randn(N, K, device=dev, dtype=torch.float64)
    Am, Bm = Mat(A, M, K), Mat(B, K, N)
    Cm = Mat.empty(M, N, dev)
    t = timeit(lambda: gemm(h, Am, Bm, Cm))
    tb = timeit(lambda: torch.matmul(B, A))          # (N,K)x(K,M): the same product in row-major storage
    ref = torch.matmul(B, A)
    err = float((Cm.buf[:N, :M] - ref).abs().max() / ref.abs().max())
    fl = 2.0 * M * N * K
    print(f"M={M} N={N} K={K}: tci_gemm {fl/t/1e12:.2f} TFLOP/s ({t*1e3:.3f} ms)  cuBLAS {fl/tb/1e12:.2f} TFLOP/s  rel.err {err:.1e}", flush=True)
