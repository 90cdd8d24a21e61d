#!/usr/bin/env python3
"""Markdown summary of an `ncu --set full` report: one column per captured launch, one row per metric.
usage: python scripts/ncu_summary.py report.ncu-rep [more.ncu-rep ...] > profiles/xyz.md   (needs `ncu` on PATH)"""
import csv
import io
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid (CTAs)"),
    ("launch__cluster_size", "cluster size"),
    ("launch__block_size", "block size"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / CTA"),
    ("sm__cycles_elapsed.max", "SM cycles"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active % (of 64 / SM)"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "FP64 pipe %"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall: barrier"),
    ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "stall: membar"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall: wait (fixed latency)"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall: short scoreboard (shared memory)"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall: long scoreboard (global / L2)"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall: math pipe throttle"),
    ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "stall: branch resolving"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared-memory bank conflicts"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
]


def short(name):
    name = name.replace("void <unnamed>::", "").replace("tci_functors::", "").replace("(int)", "")
    return name.split("(")[0]


def main():
    for rep in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units, data = rows[0], rows[1], rows[2:]
        idx = {h: i for i, h in enumerate(hdr)}
        print(f"### `{rep.split('/')[-1]}` — {len(data)} launches\n")
        print("| metric | " + " | ".join(f"#{k} `{short(r[idx['Kernel Name']])}`" for k, r in enumerate(data)) + " |")
        print("|---|" + "---|" * len(data))
        for m, label in METRICS:
            if m not in idx:
                continue
            vals = []
            for r in data:
                v = r[idx[m]]
                try:
                    x = float(v.replace(",", ""))
                    v = f"{x:,.0f}" if abs(x) >= 1000 else f"{x:.3g}"
                except ValueError:
                    pass
                vals.append(f"{v} {units[idx[m]]}".strip())
            print(f"| {label} (`{m.split('.')[0]}`) | " + " | ".join(vals) + " |")
        print()


if __name__ == "__main__":
    main()
