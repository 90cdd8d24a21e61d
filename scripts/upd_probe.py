"""Per-launch timing of the rrLU update kernel on C2-like matrices (run under gpurun); prints stats by step parity."""
import ctypes as C, sys, os, numpy as np, torch
sys.path.insert(0, '.')
import tci_b200
from tci_b200._lib import lib
from tci_b200.matrixlu import _rrlu_device, rrlu_last_timing
h = tci_b200.default_handle()
lib().tci_rrlu_last_launch_ms.restype = C.c_int64
N, R = int(os.environ.get("PN", 8192)), int(os.environ.get("PR", 512))
g = torch.Generator(device='cuda').manual_seed(1)
X = torch.randn(N, R, device='cuda', dtype=torch.float64, generator=g)
Y = torch.randn(R, N, device='cuda', dtype=torch.float64, generator=g)
sig = 2.0 ** (-torch.arange(R, device='cuda', dtype=torch.float64) / 16)
A0 = (X * sig) @ Y
for it in range(3):
    A = A0.clone()
    lib().tci_rrlu_set_profiling(h.h, 1 if it == 2 else 0)
    torch.cuda.synchronize()
    lu = _rrlu_device(h, A, N, N, N, R, 1e-14, 0.0, True)
    torch.cuda.synchronize()
    tm = rrlu_last_timing(h)
buf = (C.c_float * 4096)()
n = lib().tci_rrlu_last_launch_ms(h.h, buf, 4096)
ms = np.array(buf[:n])
upd = np.array([16.0 * (N - k - 1) ** 2 for k in range(n)])
gbs = upd / ms / 1e6
tag = os.environ.get("TAG", "")
print(f"{tag} N={N} R={R} loop_ms={tm['loop_ms']:.2f} sum_upd_ms={ms.sum():.2f} mean GB/s even={gbs[0:-1:2].mean():.0f} odd={gbs[1:-1:2].mean():.0f} all={upd[:-1].sum()/ms[:-1].sum()/1e6:.0f} "
      f"first8us={' '.join('%.0f' % (1e3 * x) for x in ms[:8])} last8us={' '.join('%.0f' % (1e3 * x) for x in ms[-9:-1])}", flush=True)
