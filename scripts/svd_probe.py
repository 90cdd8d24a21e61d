"""tci_svd_f64 (one-sided Jacobi, csrc/svd.cu) vs cuSOLVER (torch.linalg.svd) — time, sweeps, accuracy; run under gpurun."""
import sys, time, ctypes as C, numpy as np, torch
sys.path.insert(0, '.')
import tci_b200 as T
from tci_b200.dense import Mat
from tci_b200._lib import lib
h = T.default_handle(); dev = "cuda:0"
def run(m, n, reps=3):
    A = torch.randn(n, m, device=dev, dtype=torch.float64)       # Mat storage of an m x n matrix
    k = min(m, n)
    ts = []; sw = C.c_int()
    for _ in range(reps + 1):
        Am = Mat(A.clone(), m, n)
        U, Vt = Mat.empty(m, k, dev), Mat.empty(k, n, dev); S = np.zeros(k)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        h.check(lib().tci_svd_f64(h.h, Am.ptr, C.c_int64(m), C.c_int64(n), C.c_int64(Am.ld), U.ptr, C.c_int64(U.ld),
                                  S.ctypes.data_as(C.POINTER(C.c_double)), Vt.ptr, C.c_int64(Vt.ld), C.byref(sw)))
        torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    t = min(ts[1:])
    Alog = A.T
    tc = []
    for _ in range(reps + 1):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        Uc, Sc, Vc = torch.linalg.svd(Alog, full_matrices=False)
        torch.cuda.synchronize(); tc.append(time.perf_counter() - t0)
    err = float(np.max(np.abs(S - Sc.cpu().numpy())) / Sc[0].item())
    rec = float((U.view() @ torch.diag(torch.from_numpy(S).to(dev)) @ Vt.view() - Alog).abs().max() / Sc[0].item())
    wm, wn = (m, n) if m >= n else (n, m)
    rounds = sw.value * (wn - 1 + (wn & 1))
    bytes_ = rounds * 16.0 * (wm * wn + wn * wn)                 # W and V read + written once per round
    print(f"{m}x{n}: jacobi {t*1e3:8.2f} ms ({sw.value} sweeps, {rounds} launches, {bytes_/t/1e9:7.1f} GB/s algorithmic)  "
          f"cuSOLVER {min(tc[1:])*1e3:8.2f} ms   |dS|/S1 {err:.1e}  |USV-A|/S1 {rec:.1e}", flush=True)
for shape in [(800, 200), (200, 800), (2048, 512), (5120, 256), (1024, 1024), (3200, 800)]:
    run(*shape)
