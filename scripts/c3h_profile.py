"""cProfile + per-phase device-synchronised timing of the C3H sweep (full-rank hash functor, d=20, 10 sites, maxbonddim 256)."""
import cProfile, pstats, sys, time, io
sys.path.insert(0, '.')
import torch
import tci_b200 as T
ld = [20] * 10
f = T.HashFunction(ld, 1.0, 1.0)
kw = dict(tolerance=1e-10, maxbonddim=256, maxiter=2, nsearchglobalpivot=0)
T.crossinterpolate2(f, ld, **kw); torch.cuda.synchronize()
t0 = time.perf_counter(); tci, r, e = T.crossinterpolate2(f, ld, **kw); torch.cuda.synchronize(); print("wall", time.perf_counter() - t0, tci.stats)
pr = cProfile.Profile(); pr.enable()
tci, r, e = T.crossinterpolate2(f, ld, **kw); torch.cuda.synchronize()
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45); print(s.getvalue()[:9000])
