"""Pipelined (cp.async, 128x64 tile) vs single-stage DMMA GEMM vs cuBLAS; all operand orientations; run under gpurun.
TCI_GEMM_PIPE_MIN=-1 in the environment disables the pipelined kernel (read once per process), so this script is run twice."""
import os, sys, numpy as np, torch
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
print("TCI_GEMM_PIPE_MIN =", os.environ.get("TCI_GEMM_PIPE_MIN"))
for (M, N, K) in [(4096, 4096, 4096), (8192, 8192, 2048), (40000, 800, 200), (3200, 3200, 3200), (1000, 1000, 1000), (513, 515, 517), (256, 5120, 256)]:
    for ta in (False, True):
        for tb in (False, True):
            A = torch.randn((M, K) if ta else (K, M), device=dev, dtype=torch.float64)      # Mat storage: (cols, ld)
            B = torch.randn((K, N) if tb else (N, K), device=dev, dtype=torch.float64)
            Am = Mat(A, K, M) if ta else Mat(A, M, K)
            Bm = Mat(B, N, K) if tb else Mat(B, K, N)
            Cm = Mat.empty(M, N, dev)
            t = timeit(lambda: gemm(h, Am, Bm, Cm, transa=ta, transb=tb))
            a_log = A if ta else A.T                      # logical op(A): M x K
            b_log = B if tb else B.T                      # logical op(B): K x N
            ref = a_log @ b_log
            tcub = timeit(lambda: torch.matmul(a_log, b_log))
            err = float((Cm.buf[:N, :M].T - ref).abs().max() / ref.abs().max())
            fl = 2.0 * M * N * K
            print(f"M={M} N={N} K={K} ta={int(ta)} tb={int(tb)}: tci {fl/t/1e12:6.2f} TF ({t*1e3:8.3f} ms)  cuBLAS {fl/tcub/1e12:6.2f} TF  err {err:.1e}", flush=True)
            if M * N * K > 3e10 and (ta or tb):
                pass
