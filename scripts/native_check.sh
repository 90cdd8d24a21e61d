#!/bin/bash
# One GPU-box pass over the compiled sweep driver: parity tests, the same with the single-synchronisation rrLU finish
# switched off (TCI_RRLU_LATE_STATE=0; it is on by default) over the rrLU / TCI2 / dense suites, and the sweep benchmark with both engines.
export PYTHONUNBUFFERED=1
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_native_driver.py -q > gpurun_out/t_native.log 2>&1; echo "rc=$?" >> gpurun_out/t_native.log
tail -5 gpurun_out/t_native.log
timeout 150 python benchmarks/sweep_bench.py --configs c1,c3,c4 --cpu --engine both > gpurun_out/sweep_native.jsonl 2> gpurun_out/sweep_native.err; echo "rc=$?" >> gpurun_out/sweep_native.err
TCI_RRLU_LATE_STATE=0 TCI_RRLU_SMEM2=0 timeout 240 python -m pytest tests/test_gpu_native_driver.py tests/test_gpu_rrlu.py tests/test_gpu_tci2.py tests/test_gpu_dense.py tests/test_gpu_rook.py -q > gpurun_out/t_late.log 2>&1; echo "rc=$?" >> gpurun_out/t_late.log
tail -5 gpurun_out/t_late.log
TCI_RRLU_LATE_STATE=0 TCI_RRLU_SMEM2=0 timeout 120 python benchmarks/sweep_bench.py --configs c1,c3,c4 --engine both > gpurun_out/sweep_native_late.jsonl 2> gpurun_out/sweep_native_late.err; echo "rc=$?" >> gpurun_out/sweep_native_late.err
python - <<'PY'
import json
for fn in ("gpurun_out/sweep_native.jsonl", "gpurun_out/sweep_native_late.jsonl"):
    print(fn)
    for line in open(fn):
        d = json.loads(line)
        print(" ", d["config"][:12], "python %.1f ms" % (1e3 * d["gpu"]["wall_s"]), "native %.1f ms" % (1e3 * d["gpu_native"]["wall_s"]),
              "cpu %.1f ms" % (1e3 * d["cpu"]["wall_s"]) if "cpu" in d else "", "identical", d.get("native_identical_to_python"), d.get("parity"))
PY
tail -n 3 gpurun_out/sweep_native.err; tail -n 3 gpurun_out/sweep_native_late.err
(time timeout 900 python -m pytest tests -m gpu -q -x) > gpurun_out/t_all.log 2>&1; echo "rc=$?" >> gpurun_out/t_all.log
tail -n 6 gpurun_out/t_all.log
