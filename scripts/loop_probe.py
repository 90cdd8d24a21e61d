"""Whole-loop time of the device rrlu (no per-launch profiling) for several sizes; run under gpurun."""
import sys, os, numpy as np, torch
sys.path.insert(0, '.')
import tci_b200
from tci_b200.matrixlu import _rrlu_device, rrlu_last_timing
h = tci_b200.default_handle()
tag = os.environ.get("TAG", "")
for N, R in [(8192, 512), (5120, 256)] if os.environ.get("BIG") else [(8192, 512), (5120, 256), (2000, 512), (1024, 512), (512, 256), (200, 100), (100, 50)]:
    g = torch.Generator(device='cuda').manual_seed(1)
    X = torch.randn(N, R, device='cuda', dtype=torch.float64, generator=g)
    Y = torch.randn(R, N, device='cuda', dtype=torch.float64, generator=g)
    sig = 2.0 ** (-torch.arange(R, device='cuda', dtype=torch.float64) / 16)
    A0 = (X * sig) @ Y
    best = 1e9
    for it in range(4):
        A = A0.clone()
        torch.cuda.synchronize()
        lu = _rrlu_device(h, A, N, N, N, R, 1e-14, 0.0, True)
        tm = rrlu_last_timing(h)
        best = min(best, tm['loop_ms'])
    upd = sum(16.0 * (N - k - 1) ** 2 for k in range(lu.npivot - 1))
    print(f"{tag} N={N} R={R} npivot={lu.npivot} loop_ms={best:.3f} us/step={1e3*best/lu.npivot:.2f} update GB/s={upd/best/1e6:.0f}", flush=True)
