"""K1 on the C3 worst-case shape: 5120 x 5120 Pi fill (26.2 M evaluations, 209.7 MB written), Lorentzian and hash functors:
device time of the whole tci_batcheval_f64 call (CUDA events) and bit-equality with the scalar reference path (odd stride
forces the non-vectorised kernel)."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import tci_b200 as T
h = T.default_handle(); dev = "cuda:0"
ld = [20] * 20
rng = np.random.default_rng(0)
left = rng.integers(1, 21, size=(5120, 10)).astype(np.int32); right = rng.integers(1, 21, size=(5120, 10)).astype(np.int32)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, f in (("lorentz", T.LorentzFunction(ld)), ("hash", T.HashFunction(ld, 1.0, 1.0))):
    buf = torch.empty(5120 * 5120, dtype=torch.float64, device=dev)
    ts = []
    for it in range(6):
        torch.cuda.synchronize(); e0.record()
        f.evaluate_batch(left, right, 0, out=buf, stride_c=5120, stride_j=5120)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ref = torch.empty(5120 * 5121, dtype=torch.float64, device=dev)
    f.evaluate_batch(left, right, 0, out=ref, stride_c=5120, stride_j=5121)          # odd stride: scalar kernel
    same = bool(torch.equal(buf.view(5120, 5120), ref.view(5120, 5121)[:, :5120]))
    mx_v = f.evaluate_batch(left, right, 0, out=buf, stride_c=5120, stride_j=5120, maxabs=0.0)[1]
    mx_s = f.evaluate_batch(left, right, 0, out=ref, stride_c=5120, stride_j=5121, maxabs=0.0)[1]
    t = min(ts[1:])
    print(f"{name}: whole call {t:.1f} us -> {8*5120*5120/t/1e3:.0f} GB/s written, {5120*5120/t/1e3:.1f} Gevals/s; identical to scalar kernel: {same}, maxabs equal: {mx_v == mx_s}")
# write-only ceiling of this GPU for the same 209.7 MB: torch fill_ (a plain store kernel) and cudaMemset
for name, fn in (("fill_", lambda: buf.fill_(1.5)), ("zero_ (memset)", lambda: buf.zero_())):
    ts = []
    for it in range(6):
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    t = min(ts[1:])
    print(f"write-only reference {name}: {t:.1f} us -> {8*5120*5120/t/1e3:.0f} GB/s")
big = torch.empty(1 << 27, dtype=torch.float64, device=dev)          # 1 GiB
for it in range(4):
    torch.cuda.synchronize(); e0.record(); big.fill_(2.5); e1.record(); torch.cuda.synchronize()
print(f"write-only reference fill_ 1 GiB: {e0.elapsed_time(e1)*1e3:.1f} us -> {8*(1<<27)/e0.elapsed_time(e1)/1e6:.0f} GB/s")
