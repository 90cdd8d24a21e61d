"""cProfile of the latency-bound sweeps (C1: 8-d Lorentzian, C4: quantics R=40): where does the host spend a bond update?"""
import cProfile, pstats, sys, time, io, os
sys.path.insert(0, '.')
import torch
import tci_b200 as T
which = os.environ.get("CFG", "c4")
if which == "c4":
    ld = [2] * 40; f = T.QuanticsExpSin(ld); kw = dict(tolerance=1e-10, maxbonddim=512, nsearchglobalpivot=0)
    piv = [T.optfirstpivot(f, ld)]
else:
    ld = [10] * 8; f = T.LorentzFunction(ld); kw = dict(tolerance=1e-8, nsearchglobalpivot=0); piv = None
def run():
    return T.crossinterpolate2(f, ld, piv, **kw) if piv else T.crossinterpolate2(f, ld, **kw)
run(); torch.cuda.synchronize()
h = T.default_handle()
for _ in range(3):
    l0 = h.launch_count(); t0 = time.perf_counter(); tci, r, e = run(); torch.cuda.synchronize()
    print("wall", time.perf_counter() - t0, "iters", len(r), "launches", h.launch_count() - l0, tci.stats)
pr = cProfile.Profile(); pr.enable()
run(); torch.cuda.synchronize()
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(40); print(s.getvalue()[:8000])
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(40); print(s.getvalue()[:8000])
