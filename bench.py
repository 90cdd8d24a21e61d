#!/usr/bin/env python3
"""bench.py — headline benchmark of the B200-native TCI2 hot path.

Workload (BASELINE.json configs[1], "C2"): benchmark/rrlu.jl-style full-pivot rrlu on a synthetic
8192 x 8192 FP64 matrix of numerical rank 512 (A = X diag(2^{-k/16}) Y^T, X, Y ~ N(0,1) from
numpy default_rng(20260001)), `rrlu(A; maxrank=512, reltol=1e-14, abstol=0, leftorthogonal=true)`.
One step = one complete rrlu (copy of A as the reference's rrlu does, 512 pivot searches, swaps, Schur
updates, L/U extraction).

Metric: rrLU Schur-update HBM GB/s = algorithmic update bytes (16 B per updated element, SURVEY.md 8d:
516.0 GB for C2) / time of the WHOLE rrlu step.

  --gpus 1   A resident in HBM, the tuned single-GPU path (`value`); `e2e` goes through the host-buffer C-ABI call
             (`tci_rrlu_host_f64`: pinned host A in, L/U/permutations out, copies inside the timed region).
  --gpus N   STRONG scaling of the same workload: ONE C2 matrix, its columns dealt to the N ranks, factorised by the
             column-sharded rrLU (csrc/rrlu.cu tci_rrlu_dist_*: per pivot step the ranks exchange 48-byte candidate
             headers and the winner's column through peer-mapped mailboxes over NVLink, inside the step kernels).
             `value` = 516.0 GB / max-over-ranks device time; the run asserts that permutations, pivot errors and L equal
             the single-GPU factorisation's (so the scaling run doubles as the multi-GPU parity proof).

The line also carries a `sweep` object — the other half of BASELINE.json's metric: TCI2 sweep wall time and f-evals/s on
C1 / C3 / C4 (N = 1, with the oracle's CPU time beside each) and on the saturating C3H sweep (column-sharded over the N
ranks in the reference's sequential bond order).

`--impl reference` times the reference algorithm's CPU restatement (oracle/, the Julia package cannot run in this image) on
the same workload with all host threads.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 20260001
METRIC = "rrlu_schur_update_hbm_gbs"


def make_factors(n, r):
    rng = np.random.default_rng(SEED)
    X = rng.standard_normal((n, r))
    Y = rng.standard_normal((n, r))
    sig = 2.0 ** (-np.arange(r) / 16.0)
    return X, Y, sig


def update_bytes(m, n, npivot, skip_last=False):
    """Algorithmic bytes of the Schur updates of `npivot` pivot steps: 16 B x (m-k)(n-k), k = 1..npivot
    (matrixlu.jl:316-320 touches that block at every step)."""
    last = npivot - (1 if skip_last else 0)
    return float(sum(16.0 * (m - k) * (n - k) for k in range(1, last + 1)))


class ClockSampler(threading.Thread):
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.rows, self._stop_ev, self.proc = gpu_index, [], threading.Event(), None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
                if self._stop_ev.is_set():
                    break
        except Exception:
            pass

    def stop(self):
        self._stop_ev.set()
        if self.proc:
            self.proc.terminate()

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def kernel_source_sha():
    p = os.path.join(ROOT, "tensorcrossinterpolation.jl_b200", "csrc", "rrlu.cu")
    return hashlib.sha256(open(p, "rb").read()).hexdigest()[:16]


def measured_traffic():
    """DRAM bytes per launch of the update kernel from the committed `ncu --set full` capture — only when that capture was
    taken on THIS version of csrc/rrlu.cu (the json records the source hash); a stale capture is reported as such, loudly."""
    tp = os.path.join(ROOT, "profiles", "traffic_update_kernel.json")
    if not os.path.exists(tp):
        return None, "no ncu capture committed"
    t = json.load(open(tp))
    if t.get("rrlu_cu_sha16") != kernel_source_sha():
        print(f"bench.py: WARNING: profiles/traffic_update_kernel.json was captured on another version of csrc/rrlu.cu "
              f"({t.get('rrlu_cu_sha16')} != {kernel_source_sha()}): roofline.traffic is reported as null", file=sys.stderr)
        return None, f"stale: ncu capture {t.get('source')} predates the current csrc/rrlu.cu"
    return t.get("dram_bytes_per_launch"), f"{t.get('source')} (rrlu.cu {t.get('rrlu_cu_sha16')})"


def cpu_rrlu(A_host, steps, threads):
    """The oracle's rrlu loop (reference algorithm, matrixlu.jl:346-396) on the first `steps` pivots of the matrix with
    `threads` host threads (1 = the scalar loop, the reference's own execution model)."""
    from oracle import oracle as O
    m, n = A_host.shape
    t0 = time.perf_counter()
    lu = O.rrlu(A_host, maxrank=steps, threads=threads)
    dt = time.perf_counter() - t0
    return update_bytes(m, n, lu.npivot) / dt / 1e9, dt, lu


def host_threads():
    """All host cores this process may use (torchrun exports OMP_NUM_THREADS=1; the oracle's loops take an explicit thread
    count, so that setting does not apply)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def run_reference(args):
    """Reference arm: the oracle port of the reference's rrlu on the host cores (all threads), same matrix, same call.
    A step is the WHOLE 512-pivot factorisation when --steps of them fit the budget (they do on the 16-core box: ~6 s
    each), else the first s pivots, s sized from the warm-up rate."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    n, r = args.size, args.rank
    X, Y, sig = make_factors(n, r)
    A = np.asfortranarray((X * sig) @ Y.T)
    nt = host_threads()
    probe = min(32, r)
    rate = None
    for _ in range(max(1, args.warmup)):
        gbs, dt, _ = cpu_rrlu(A, probe, nt)
        rate = gbs if rate is None else max(rate, gbs)
    est_full = update_bytes(n, n, r) / 1e9 / rate
    sample = r
    if est_full * args.steps > args.reference_budget_s:
        target = args.reference_budget_s / args.steps * rate * 1e9
        sample = probe
        while sample < r and update_bytes(n, n, sample + 1) <= target:
            sample += 1
    t_tot, b_tot = 0.0, 0.0
    for _ in range(args.steps):
        gbs, dt, lu = cpu_rrlu(A, sample, nt)
        t_tot += dt; b_tot += gbs * dt
    val = b_tot / t_tot
    what = "the whole factorisation" if sample == r else f"the first {sample} of {r} pivot steps"
    desc = f"{what} of the {n}x{n} rrlu per step (oracle/tci_oracle.cpp, {nt} threads)"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps,
            "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"C2: rrlu {n}x{n} FP64, numerical rank {r}, maxrank={r}, reltol=1e-14", "sample": desc,
                       "npivot": int(lu.npivot), "full_workload": sample == r},
            "cpu_baseline": {"value": val, "unit": "GB/s", "cores": nt, "kind": "port", "sample": desc},
            "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---- the sweep half of the metric ------------------------------------------------------------------------------------
def _sets_equal(t, o):
    return bool(all(np.array_equal(t.Iset[b], o.Iset(b + 1)) and np.array_equal(t.Jset[b], o.Jset(b + 1))
                    for b in range(len(t.localdims))))


def _timed_sweep(T, torch, h, make, ld, reps, pivots=None, **kw):
    best = None
    for _ in range(reps + 1):                              # the first run warms allocator / workspaces
        f = make()
        torch.cuda.synchronize()
        l0, t0 = h.launch_count(), time.perf_counter()
        tci, ranks, errors = T.crossinterpolate2(f, ld, pivots, nsearchglobalpivot=0, **kw)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        rec = {"wall_ms": 1e3 * dt, "rank": int(ranks[-1]), "iterations": len(ranks), "half_sweeps": 2 * len(ranks),
               "ms_per_half_sweep": 1e3 * dt / (2 * len(ranks)), "evals": int(tci.stats["nevals"]),
               "evals_per_s": tci.stats["nevals"] / dt, "final_error": float(errors[-1]), "gpu_launches": h.launch_count() - l0}
        if best is None or rec["wall_ms"] < best[0]["wall_ms"]:
            best = (rec, tci)
    return best


def _cpu_sweep(O, kind, params, ld, pivots=None, **kw):
    f = O.Func(kind, params)
    t0 = time.perf_counter()
    tci, ranks, errors = O.crossinterpolate2(f, ld, pivots, **kw)
    dt = time.perf_counter() - t0
    return {"wall_ms": 1e3 * dt, "rank": int(ranks[-1]), "evals": int(tci.stats()["nevals"]), "cores": 1,
            "kind": "port (oracle/tci_oracle.cpp, scalar loops like the reference)"}, tci


def sweep_single_gpu(T, torch, h, engine):
    """C1 / C3 / C4 (BASELINE configs 0, 2, 3) and the saturating C3H sweep on ONE GPU, the oracle on the host beside each."""
    from oracle import oracle as O
    out = {"engine": engine}
    ld = [10] * 8
    g, gt = _timed_sweep(T, torch, h, lambda: T.LorentzFunction(ld, handle=h), ld, 3, tolerance=1e-8, engine=engine, handle=h)
    c, ct = _cpu_sweep(O, O.F_LORENTZ, [1.0], ld, tolerance=1e-8)
    out["C1"] = {"config": "8-d Lorentzian, localdims=fill(10,8), tolerance=1e-8", "gpu": g, "cpu": c,
                 "pivot_sets_identical": _sets_equal(gt, ct)}
    ld = [20] * 20
    g, gt = _timed_sweep(T, torch, h, lambda: T.LorentzFunction(ld, handle=h), ld, 2, tolerance=1e-10, maxbonddim=256, engine=engine,
                         handle=h)
    c, ct = _cpu_sweep(O, O.F_LORENTZ, [1.0], ld, tolerance=1e-10, maxbonddim=256)
    out["C3"] = {"config": "20-d Lorentzian, localdims=fill(20,20), maxbonddim=256, tolerance=1e-10", "gpu": g, "cpu": c,
                 "pivot_sets_identical": _sets_equal(gt, ct)}
    ld = [2] * 40
    f0 = T.QuanticsExpSin(ld, handle=h)
    first = [list(T.optfirstpivot(f0, ld))]
    g, gt = _timed_sweep(T, torch, h, lambda: T.QuanticsExpSin(ld, handle=h), ld, 3, pivots=first, tolerance=1e-10, maxbonddim=512,
                         engine=engine, handle=h)
    c, ct = _cpu_sweep(O, O.F_QEXPSIN, [10.0, 100.0, 1.1], ld, first, tolerance=1e-10, maxbonddim=512)
    out["C4"] = {"config": "quantics R=40, exp(-10x) sin(2 pi 100 x^1.1), maxbonddim=512, tolerance=1e-10, optfirstpivot", "gpu": g, "cpu": c,
                 "ranks_equal": g["rank"] == c["rank"],
                 "note": "libm (CPU leg) and libdevice differ in the last ulp; bit-exact index-set parity against the oracle fed the "
                         "device values is tests/test_gpu_baseline_configs.py::test_c4_full_size_device_fed_oracle"}
    ld = [20] * 10
    g, gt = _timed_sweep(T, torch, h, lambda: T.HashFunction(ld, 1.0, 1.0, handle=h), ld, 2, tolerance=1e-10, maxbonddim=256, maxiter=2,
                         engine=engine, handle=h)
    gs, gts = _timed_sweep(T, torch, h, lambda: T.HashFunction(ld, 1.0, 1.0, handle=h), ld, 1, tolerance=1e-10, maxbonddim=64, maxiter=1,
                           engine=engine, handle=h)
    c, ct = _cpu_sweep(O, O.F_HASH, [1.0, 1.0], ld, tolerance=1e-10, maxbonddim=64, maxiter=1)
    c["sample"] = "the same functor at maxbonddim=64, maxiter=1 (the full configuration would take minutes on one core)"
    out["C3H"] = {"config": "full-rank hash functor, localdims=fill(20,10), maxbonddim=256, maxiter=2 (Pi up to 5120 x 5120)", "gpu": g,
                  "gpu_on_cpu_sample": gs, "cpu": c, "pivot_sets_identical_on_sample": _sets_equal(gts, ct)}
    return out


def sweep_multi_gpu(T, torch, dist, h, comm, world, rank):
    """The saturating C3H sweep in the reference's sequential bond order with the columns of every large Pi spread over the
    ranks (K1 evaluates only the local column block, column-sharded rrLU, only U gathered); max over ranks; index sets
    must equal the one-GPU run's."""
    ld = [20] * 10
    kw = dict(tolerance=1e-10, maxbonddim=256, maxiter=2, nsearchglobalpivot=0, handle=h)
    f = T.HashFunction(ld, 1.0, 1.0, handle=h)
    dev = torch.device(f"cuda:{h.device}")

    def timed(**extra):
        best, tci = None, None
        for _ in range(2):
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            tci = T.crossinterpolate2(f, ld, **kw, **extra)[0]
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        t = torch.tensor([best], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), tci
    t1, ref = timed()
    tn, sh = timed(comm=comm)
    same = all(np.array_equal(a, b) for a, b in zip(ref.Iset, sh.Iset)) and all(np.array_equal(a, b) for a, b in zip(ref.Jset, sh.Jset))
    flag = torch.tensor([1.0 if same else 0.0], dtype=torch.float64, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ev = int(ref.stats["nevals"])
    extra = {}
    try:
        extra.update(_k1_sharded(T, torch, dist, h, comm, world, dev))
        extra.update(_oddeven_sweep(T, torch, dist, f, ld, kw, comm, dev, timed, t1))
    except Exception as ex:                                          # never lose the headline to an auxiliary measurement
        extra["partitions_error"] = f"{type(ex).__name__}: {ex}"
    return {**extra,
            "C3H": {"config": "full-rank hash functor, localdims=fill(20,10), maxbonddim=256, maxiter=2; sequential bond order, columns of "
                              "every Pi >= 2^20 entries spread over the ranks",
                    "one_gpu_wall_ms": 1e3 * t1, "wall_ms": 1e3 * tn, "speedup": t1 / tn, "evals": ev, "evals_per_s": ev / tn,
                    "sharded_bonds": int(sh.stats.get("nsharded", 0)), "pivot_sets_identical_to_one_gpu": bool(flag.item() == 1.0)}}


K1_HEAVY_SRC = """
__device__ double tci_user_f(const int* idx, int n, const double* p, int np) {
    double s = 0.0;
    for (int i = 0; i < n; ++i) s += sin(p[0] * idx[i] * (i + 1)) * exp(-p[1] * idx[i]);
    return s / (1.0 + s * s);
}
"""


def _k1_sharded(T, torch, dist, h, comm, world, dev):
    """north_star partition 1: ONE Pi fill of the C3 worst-case shape (5120 x 5120, 26.2 M evaluations, 210 MB) with
    the evaluations split over the ranks and every rank's block stored straight into every peer's copy of Pi by the evaluation
    kernel (fused compute + all-gather over peer memory); against the same fill done entirely on every rank.  Two functors: the
    cheap built-in hash functor (HBM-write-bound: the split cannot pay, NVLink moves 8 B per value) and an arithmetic-heavy
    NVRTC plugin functor (20 sin + 20 exp per value: the split pays).  Device-side wall time, max over ranks; the assembled
    matrix must equal the local one bit for bit on every rank."""
    ld = [20] * 10
    rng = np.random.default_rng(0)
    nI = nJ = 5120
    left = rng.integers(1, 21, size=(nI, 5)).astype(np.int32)
    right = rng.integers(1, 21, size=(nJ, 5)).astype(np.int32)
    buf = torch.empty(nI * nJ, dtype=torch.float64, device=dev)

    def timed(fn, reps=5):
        fn(); fn()
        ts = []
        for _ in range(reps):
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        t = torch.tensor([min(ts)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    out = {"config": "Pi fill 5120 x 5120 (C3 worst-case bond), columns of Pi over the ranks; per functor: everything on every rank / split + NCCL "
                     "all-gather / split + fused peer stores", "evals": nI * nJ}
    for name, f in (("hash", T.HashFunction(ld, 1.0, 1.0, handle=h)), ("heavy_plugin", T.UserFunction(K1_HEAVY_SRC, ld, [0.37, 0.05], handle=h))):
        t_local = timed(lambda: f.evaluate_batch(left, right, 0, out=buf, stride_c=nI, stride_j=nI, maxabs=0.0))
        ga = T.ShardedBatchEvaluator(f, comm)
        t_gather = timed(lambda: ga.evaluate_batch(left, right, 0, out=buf, stride_c=nI, stride_j=nI, maxabs=0.0))
        fu = T.ShardedBatchEvaluator(f, comm, fused=True)

        def fused():
            m = fu.alloc_matrix(nI, nJ, 16)
            o = m.buf if m is not None else buf
            fu.evaluate_batch(left, right, 0, out=o, stride_c=nI, stride_j=nI, maxabs=0.0)
            return o
        t_fused = timed(fused)
        ref = f.evaluate_batch(left, right, 0, out=torch.empty_like(buf), stride_c=nI, stride_j=nI)[0]
        o = fused()
        torch.cuda.synchronize(); dist.barrier()
        same = torch.tensor([1.0 if torch.equal(o.reshape(-1)[: nI * nJ], ref.reshape(-1)[: nI * nJ]) else 0.0], dtype=torch.float64, device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        out[name] = {"local_ms": 1e3 * t_local, "sharded_nccl_allgather_ms": 1e3 * t_gather, "sharded_fused_peer_store_ms": 1e3 * t_fused,
                     "evals_per_s": nI * nJ / min(t_local, t_fused), "speedup_fused_vs_local": t_local / t_fused,
                     "nvlink_GBps_per_rank": 8 * nI * nJ * (world - 1) / world / t_fused / 1e9,
                     "identical_to_local_on_every_rank": bool(same.item() == 1.0)}
        del fu, ga
    return {"K1_sharded": out}


def _oddeven_sweep(T, torch, dist, f, ld, kw, comm, dev, timed, t_seq):
    """north_star partition 2: the parallel-bond sweep — bonds of one parity are independent, so they are dealt to
    the ranks and the new pivot sets exchanged (NCCL) after each parity.  The schedule differs from the reference's sequential
    one by construction (tests compare it with the oracle's odd/even schedule); here: index sets of the N-rank run must equal
    the same schedule run on one GPU."""
    t_oe1, one = timed(bondorder="oddeven")
    t_oen, par = timed(bondorder="oddeven", comm=comm)
    same = all(np.array_equal(a, b) for a, b in zip(one.Iset, par.Iset)) and all(np.array_equal(a, b) for a, b in zip(one.Jset, par.Jset))
    flag = torch.tensor([1.0 if same else 0.0], dtype=torch.float64, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    evt = torch.tensor([float(par.stats["nevals"])], dtype=torch.float64, device=dev)      # every rank evaluates only its own bonds
    dist.all_reduce(evt, op=dist.ReduceOp.SUM)
    ev = int(evt.item())
    return {"C3H_parallel_bonds": {"config": "the C3H sweep with bondorder='oddeven': bonds of one parity spread over the ranks, pivot sets exchanged per parity",
                                   "sequential_one_gpu_wall_ms": 1e3 * t_seq, "oddeven_one_gpu_wall_ms": 1e3 * t_oe1, "wall_ms": 1e3 * t_oen,
                                   "speedup_vs_sequential": t_seq / t_oen, "evals": ev, "evals_per_s": ev / t_oen,
                                   "pivot_sets_identical_to_one_gpu_same_schedule": bool(flag.item() == 1.0)}}


# ---- the rrLU headline -----------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    import tci_b200
    from tci_b200._lib import lib, p_f64, p_i64
    from tci_b200.matrixlu import _rrlu_device, rrlu_last_timing

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL prints its version banner from C to file descriptor 1.  Keep stdout to the ONE JSON line: Python's sys.stdout moves
        # to a duplicate of the original descriptor, descriptor 1 itself is pointed at stderr for everything native.
        sys.stdout.flush()
        keep = os.dup(1)
        os.dup2(2, 1)
        sys.stdout = os.fdopen(keep, "w", buffering=1)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    dev = torch.device(f"cuda:{local}")
    h = tci_b200.Handle(local)
    h.use_torch_stream()

    n, r = args.size, args.rank
    X, Y, sig = make_factors(n, r)
    Xd = torch.from_numpy(X * sig).to(dev)
    Yd = torch.from_numpy(Y).to(dev)
    A0 = (Yd @ Xd.T).contiguous()          # stored (n, lda=m): row j = column j of the logical A = X diag(sig) Y^T
    del Xd, Yd
    peak, peak_src = peaks()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if world > 1:
        return run_sharded(args, tci_b200, torch, dist, h, dev, A0, world, rank, local, peak, peak_src, barrier, max_over_ranks)

    work = torch.empty_like(A0)

    def step_device():
        work.copy_(A0)                       # rrlu copies A (matrixlu.jl:462)
        return _rrlu_device(h, work, n, n, n, r, 1e-14, 0.0, True, extract=True)

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        lu = step_device()
    npivot = lu.npivot
    barrier()
    l0 = h.launch_count()
    e0.record()
    for _ in range(args.steps):
        lu = step_device()
    e1.record()
    barrier()
    launches = h.launch_count() - l0
    ms_per_step = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    alg = update_bytes(n, n, npivot)
    value = alg / (ms_per_step * 1e-3) / 1e9

    # ---- e2e: host buffers through the C ABI ------------------------------------------------------
    A_host_t = torch.empty((n, n), dtype=torch.float64).pin_memory()
    A_host_t.copy_(A0)                       # same (n, lda) storage == column-major A
    A_host = A_host_t.numpy()
    Lh = torch.empty((r, n), dtype=torch.float64).pin_memory().numpy()      # column-major m x r
    Uh = torch.empty((n, r), dtype=torch.float64).pin_memory().numpy()      # column-major r x n
    rowperm = np.zeros(n, dtype=np.int64); colperm = np.zeros(n, dtype=np.int64); pe = np.zeros(r + 1)
    npv, err = C.c_int64(), C.c_double()

    def step_host():
        h.check(lib().tci_rrlu_host_f64(h.h, p_f64(A_host), C.c_int64(n), C.c_int64(n), C.c_int64(n), C.c_int64(r),
                                        C.c_double(1e-14), C.c_double(0.0), 1, p_i64(rowperm), p_i64(colperm),
                                        C.byref(npv), C.byref(err), p_f64(Lh), C.c_int64(n), p_f64(Uh), C.c_int64(r), p_f64(pe)))

    for _ in range(min(args.warmup, 2)):
        step_host()
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_host()
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    e2e_val = alg / (e2e_ms * 1e-3) / 1e9
    sampler.stop()
    h2d = 8 * n * n
    d2h = 8 * (n * npivot + npivot * n) + 8 * 2 * n + 8 * (npivot + 1)
    assert npv.value == npivot and np.array_equal(rowperm, lu.rowpermutation)

    # ---- roofline of the dominant kernel (rrlu_update_kernel): per-launch event pairs on a separate pass
    lib().tci_rrlu_set_profiling(h.h, 1)
    step_device()
    tm = rrlu_last_timing(h)
    lib().tci_rrlu_set_profiling(h.h, 0)
    upd_ms = float(lib_ms1(h))
    alg_kernel = update_bytes(n, n, npivot, skip_last=True)          # the last pivot's update is never needed
    achieved = alg_kernel / (upd_ms * 1e-3) / 1e9 if upd_ms > 0 else None
    traffic, traffic_src = measured_traffic()
    roof = {"bound": "hbm", "kernel": "rrlu_update_kernel<UPDATE,VEC>", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": (achieved / peak) if achieved else None, "traffic": traffic, "traffic_source": traffic_src,
            "peak_source": peak_src, "launches": int(tm["steps"]), "avg_launch_us": 1e3 * upd_ms / max(1, tm["steps"]),
            "algorithmic_bytes_per_launch_avg": alg_kernel / max(1, tm["steps"] - 1),
            "kernel_share_of_loop": upd_ms / tm["loop_ms"] if tm["loop_ms"] else None}

    # ---- CPU baseline on a bounded sample, and its pivots against the GPU's ---------------------------------------
    nt = host_threads()
    ks = min(args.cpu_sample_steps, npivot)
    gbs, dt, ref = cpu_rrlu(np.asfortranarray(A_host.T), ks, nt)
    same = (np.array_equal(ref.rowpermutation[:ks], lu.rowpermutation[:ks]) and np.array_equal(ref.colpermutation[:ks], lu.colpermutation[:ks])
            and np.array_equal(ref.pivoterrors()[:ks], lu.pivoterrors()[:ks]))
    assert same, "the GPU's first pivots differ from the oracle's"
    cpu = {"value": gbs, "unit": "GB/s", "cores": nt, "kind": "port",
           "sample": f"first {ref.npivot} of {npivot} pivot steps of the same {n}x{n} matrix, {dt:.1f} s, oracle/tci_oracle.cpp on {nt} threads",
           "pivots_equal_gpu": bool(same)}

    sweep = None
    if not args.no_sweep:
        h.use_torch_stream()
        sweep = sweep_single_gpu(tci_b200, torch, h, args.engine)

    line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"C2: rrlu {n}x{n} FP64, numerical rank {r}, maxrank={r}, reltol=1e-14",
                       "npivot": npivot, "parallelism": "single GPU",
                       "l2": f"input {8 * n * n / 1e6:.0f} MB > L2 (126 MB): no flush needed",
                       "algorithmic_update_GB_per_step": alg / 1e9},
            "e2e": {"value": e2e_val, "unit": "GB/s", "ms_per_step": e2e_ms, "steps": args.steps, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "api": "tci_rrlu_host_f64 (pinned host A in; L, U, perms out)"},
            "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "clocks": sampler.summary(), "sweep": sweep}
    print(json.dumps(line), flush=True)


def run_sharded(args, T, torch, dist, h, dev, A0, world, rank, local, peak, peak_src, barrier, max_over_ranks):
    """--gpus N > 1: one C2 matrix, columns over the ranks, fused mailbox exchange (strong scaling)."""
    from tci_b200.matrixlu import _rrlu_device
    from tci_b200.parallel import PeerBuffers, ShardedRRLU, partition, rrlu_sharded
    n, r = args.size, args.rank
    comm = T.Comm()
    s, e = partition(n, world)[rank]
    nloc = e - s
    blk0 = A0[s:e].contiguous()                     # (nloc, m) row-major == column-major m x nloc: this rank's columns, resident
    probe = ShardedRRLU(T.Mat(blk0.clone(), n, nloc), n, s, rank, world, r, handle=h)
    peers = PeerBuffers(h, comm, probe.mailbox_doubles(), nslots=1)
    probe.close()
    work = torch.empty_like(blk0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def step_device():
        work.copy_(blk0)                            # rrlu copies A (matrixlu.jl:462)
        return rrlu_sharded(T.Mat(work, n, nloc), n, s, comm, maxrank=r, handle=h, check_every=0, fused=True, peers=peers)

    sampler = ClockSampler(local)                   # started before the warm-up: nvidia-smi needs a moment on an 8-GPU box
    sampler.start()
    for _ in range(args.warmup):
        out = step_device()
    npivot = int(out["npivot"])
    barrier()
    l0 = h.launch_count()
    e0.record()
    for _ in range(args.steps):
        out = step_device()
    e1.record()
    barrier()
    launches = h.launch_count() - l0
    ms_per_step = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    alg = update_bytes(n, n, npivot)
    value = alg / (ms_per_step * 1e-3) / 1e9

    # ---- e2e: this rank's column block from pinned host memory, its factors back ---------------------------------
    blk_host = torch.empty(blk0.shape, dtype=torch.float64).pin_memory()
    blk_host.copy_(blk0)
    L_host = torch.empty((r, n), dtype=torch.float64).pin_memory()          # Mat storage is (cols, ld): row k = column k of L
    U_host = torch.empty((nloc, r), dtype=torch.float64).pin_memory()       # row j = local column j of U

    def step_host():
        work.copy_(blk_host, non_blocking=True)
        o = rrlu_sharded(T.Mat(work, n, nloc), n, s, comm, maxrank=r, handle=h, check_every=0, fused=True, peers=peers)
        k = o["npivot"]
        U_host[:, :k].copy_(o["Uloc"].buf[:nloc, :k], non_blocking=True)
        if rank == 0:
            L_host[:k].copy_(o["L"].buf[:k, :n], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return o

    step_host()
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_host()
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    sampler.stop()
    h2d = 8 * n * n                                  # all ranks together: every column is uploaded once
    d2h = 8 * (n * npivot + npivot * n) + 8 * 2 * n + 8 * (npivot + 1)

    # ---- parity with the single-GPU factorisation of the same matrix (rank 0 runs it; every rank holds the same sharded result)
    ok = 1.0
    if rank == 0:
        full = A0.clone()
        ref = _rrlu_device(h, full, n, n, n, r, 1e-14, 0.0, True, extract=True)
        same = (ref.npivot == npivot and np.array_equal(ref.rowpermutation, out["rowperm"]) and np.array_equal(ref.colpermutation, out["colperm"])
                and np.array_equal(ref.pivoterrors(), out["pivoterrors"])
                and bool(torch.equal(ref.L[:, :npivot], out["L"].view()[:, :npivot])))
        ok = 1.0 if same else 0.0
        del full
    flag = torch.tensor([ok], dtype=torch.float64, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    parity = bool(flag.item() == 1.0)

    sweep = None
    if not args.no_sweep:
        sweep = sweep_multi_gpu(T, torch, dist, h, comm, world, rank)
    peers.close()
    if rank == 0:
        assert parity, "column-sharded rrLU differs from the single-GPU factorisation"
        per_gpu = value / world
        l2note = (" <= L2 (126 MB): the update runs from L2, so the per-GPU figure can exceed the HBM peak" if 8 * n * nloc <= 126e6
                  else " > L2 (126 MB): no flush needed")
        line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"C2: rrlu {n}x{n} FP64, numerical rank {r}, maxrank={r}, reltol=1e-14", "npivot": npivot,
                           "parallelism": f"ONE matrix, columns over {world} GPUs (column-sharded rrLU, in-kernel mailbox exchange over NVLink)",
                           "l2": f"local block {8 * n * nloc / 1e6:.0f} MB per GPU" + l2note,
                           "algorithmic_update_GB_per_step": alg / 1e9},
                "e2e": {"value": alg / (e2e_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": e2e_ms, "steps": args.steps,
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "api": "rrlu_sharded (tci_rrlu_dist_*): pinned host column blocks in; L (rank 0) and the U blocks out"},
                "gpu_launches": int(launches) * world,
                "roofline": {"bound": "hbm", "kernel": "rrlu_dist_update_kernel (whole step, per GPU)", "achieved": per_gpu, "peak": peak,
                             "unit": "GB/s", "frac": per_gpu / peak, "traffic": None, "peak_source": peak_src,
                             "note": "whole-step figure per GPU (exchange and launch gaps included), not a kernel-alone time"},
                "parity": {"identical_to_single_gpu": parity, "what": "row/column permutations, pivot errors, L"},
                "cpu_baseline": None, "clocks": sampler.summary(), "sweep": sweep}
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()


def lib_ms1(h):
    from tci_b200._lib import lib, p_f64
    ms = np.zeros(3)
    steps = C.c_int64()
    lib().tci_rrlu_last_timing(h.h, p_f64(ms), C.byref(steps))
    return ms[1]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--rank", type=int, default=512)
    ap.add_argument("--cpu-sample-steps", type=int, default=128)
    ap.add_argument("--reference-budget-s", type=float, default=240.0)
    ap.add_argument("--no-sweep", action="store_true", help="skip the `sweep` object (C1/C3/C4/C3H sweep times)")
    ap.add_argument("--engine", default="auto", help="crossinterpolate2 engine for the sweep object")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
