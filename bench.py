#!/usr/bin/env python3
"""bench.py — headline benchmark of the B200-native TCI2 hot path.

Workload (BASELINE.json configs[1], "C2"): benchmark/rrlu.jl-style full-pivot rrlu on a synthetic
8192 x 8192 FP64 matrix of numerical rank 512 (A = X diag(2^{-k/16}) Y^T, X, Y ~ N(0,1) from
numpy default_rng(20260001)), `rrlu(A; maxrank=512, reltol=1e-14, abstol=0, leftorthogonal=true)`.
One step = one complete rrlu (copy of A as the reference's rrlu does, 512 pivot searches, swaps, Schur
updates, L/U extraction).

Metric: rrLU Schur-update HBM GB/s = algorithmic update bytes (16 B per updated element, SURVEY.md 8d:
515.9 GB for C2) / time of the WHOLE rrlu step, summed over all GPUs.  `value` has A resident in HBM;
`e2e` goes through the host-buffer C-ABI call (`tci_rrlu_host_f64`: pinned host A in, L/U/permutations
out, copies inside the timed region).  rrLU does not shard: with --gpus N every rank
factorises its own replica (weak scaling, no collective on the data path).

`--impl reference` times the reference algorithm's CPU restatement (oracle/, the Julia package cannot
run in this image) on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
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


def cpu_rrlu_sample(A_host, sample_steps):
    """The oracle's rrlu (reference algorithm, single thread like the Julia loops matrixlu.jl:316-320) on
    the first `sample_steps` pivots of the same matrix."""
    from oracle import oracle as O
    m, n = A_host.shape
    t0 = time.perf_counter()
    lu = O.rrlu(A_host, maxrank=sample_steps)
    dt = time.perf_counter() - t0
    # the reference runs a search + update per pivot and copies A first; count the same update bytes
    return update_bytes(m, n, lu.npivot) / dt / 1e9, dt, lu.npivot


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n, r = args.size, args.rank
    X, Y, sig = make_factors(n, r)
    A = np.asfortranarray((X * sig) @ Y.T)
    sample = args.cpu_sample_steps
    for _ in range(args.warmup):
        cpu_rrlu_sample(A, 2)
    t_tot, b_tot = 0.0, 0.0
    for _ in range(args.steps):
        gbs, dt, npiv = cpu_rrlu_sample(A, sample)
        t_tot += dt; b_tot += gbs * dt
    val = b_tot / t_tot
    desc = f"first {sample} of {r} pivot steps of the {n}x{n} rrlu per step (oracle/tci_oracle.cpp, 1 thread)"
    line = {"impl": "reference", "metric": "rrlu_schur_update_hbm_gbs", "value": val, "unit": "GB/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"C2: rrlu {n}x{n} FP64, numerical rank {r}, maxrank={r}", "sample": desc},
            "cpu_baseline": {"value": val, "unit": "GB/s", "cores": 1, "kind": "port", "sample": desc},
            "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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
    work = torch.empty_like(A0)

    def step_device():
        work.copy_(A0)                       # rrlu copies A (matrixlu.jl:462)
        return _rrlu_device(h, work, n, n, n, r, 1e-14, 0.0, True, extract=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        lu = step_device()
    npivot = lu.npivot
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    l0 = h.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        lu = step_device()
    e1.record()
    barrier()
    launches = h.launch_count() - l0
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    alg = update_bytes(n, n, npivot)
    value = world * alg / (ms_per_step * 1e-3) / 1e9

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

    step_host()
    barrier()
    ksteps = max(1, min(args.steps, 3))
    e0.record()
    for _ in range(ksteps):
        step_host()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item()) / ksteps
    e2e_val = world * alg / (e2e_ms * 1e-3) / 1e9
    sampler.stop()
    h2d = 8 * n * n
    d2h = 8 * (n * npivot + npivot * n) + 4 * 2 * n + 8 * npivot
    assert npv.value == npivot and np.array_equal(rowperm, lu.rowpermutation)

    # ---- roofline of the dominant kernel (rrlu_update_kernel): per-launch event pairs on a separate pass
    lib().tci_rrlu_set_profiling(h.h, 1)
    step_device()
    tm = rrlu_last_timing(h)
    lib().tci_rrlu_set_profiling(h.h, 0)
    upd_ms = float(lib_ms1(h))
    peak, peak_src = peaks()
    alg_kernel = update_bytes(n, n, npivot, skip_last=True)          # the last pivot's update is never needed
    achieved = alg_kernel / (upd_ms * 1e-3) / 1e9 if upd_ms > 0 else None
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic_update_kernel.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get("dram_bytes_per_launch")
    roof = {"bound": "hbm", "kernel": "rrlu_update_kernel<UPDATE,VEC>", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": (achieved / peak) if achieved else None, "traffic": traffic, "peak_source": peak_src,
            "launches": int(tm["steps"]), "avg_launch_us": 1e3 * upd_ms / max(1, tm["steps"]),
            "algorithmic_bytes_per_launch_avg": alg_kernel / max(1, tm["steps"] - 1),
            "kernel_share_of_loop": upd_ms / tm["loop_ms"] if tm["loop_ms"] else None}

    if rank == 0:
        cpu = None
        if world == 1 or True:
            gbs, dt, npv_cpu = cpu_rrlu_sample(np.asfortranarray(A_host.T), args.cpu_sample_steps)
            cpu = {"value": gbs, "unit": "GB/s", "cores": 1, "kind": "port",
                   "sample": f"first {npv_cpu} of {npivot} pivot steps of the same {n}x{n} matrix, {dt:.1f} s, oracle/tci_oracle.cpp"}
        line = {"metric": "rrlu_schur_update_hbm_gbs", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"C2: rrlu {n}x{n} FP64, numerical rank {r}, maxrank={r}, reltol=1e-14",
                           "npivot": npivot, "parallelism": "replicas only (rrLU does not shard)" if world > 1 else "single GPU",
                           "l2": f"input {8 * n * n / 1e6:.0f} MB > L2 (126 MB): no flush needed",
                           "algorithmic_update_GB_per_step": alg / 1e9},
                "e2e": {"value": e2e_val, "unit": "GB/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h, "api": "tci_rrlu_host_f64 (pinned host A in; L, U, perms out)"},
                "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "clocks": sampler.summary()}
        print(json.dumps(line), flush=True)
    if world > 1:
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
    ap.add_argument("--cpu-sample-steps", type=int, default=32)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
