import time, numpy as np, torch, sys
sys.path.insert(0, '.')
import tci_b200
from tci_b200.matrixlu import _rrlu_device, rrlu_last_timing
h = tci_b200.default_handle()
for N, R in [(8192, 512), (5120, 256), (1024, 512), (2000, 512)]:
    g = torch.Generator(device='cuda').manual_seed(1)
    X = torch.randn(N, R, device='cuda', dtype=torch.float64, generator=g)
    Y = torch.randn(R, N, device='cuda', dtype=torch.float64, generator=g)
    sig = 2.0 ** (-torch.arange(R, device='cuda', dtype=torch.float64) / 16)
    A0 = (X * sig) @ Y
    for it in range(3):
        A = A0.clone()
        torch.cuda.synchronize(); t = time.time()
        lu = _rrlu_device(h, A, N, N, N, R, 1e-14, 0.0, True)
        torch.cuda.synchronize(); dt = time.time() - t
        tm = rrlu_last_timing(h)
        upd = sum(16.0 * (N - k - 1) ** 2 for k in range(lu.npivot - 1))
        print(N, R, 'npivot', lu.npivot, 'wall %.2f ms' % (dt * 1e3), tm, 'update GB/s (loop) %.0f' % (upd / tm['loop_ms'] / 1e6), flush=True)
