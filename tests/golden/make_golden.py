#!/usr/bin/env python3
"""Extract the literal known-answer matrices from the reference's own test files into
tests/golden/reference_literals.json.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):
    python tests/golden/make_golden.py
The JSON it writes is committed; tests read the JSON, never /root/reference.

Sources:
  test/test_matrixlu.jl    :8-19 argmax 10x8, :55-60 rrlu 4x4, :89-93 rank-1 3x3, :100-109 8x6,
                           :142-158 p (10x3) and q (3x10), :168-171 identity, :178-184 5x5, :198-203 1e-13 * 4x4
  test/test_matrixluci.jl  :7-16 8x6 (differs from matrixlu's in one entry), :49-66 p, q
"""
import json
import os
import re

REF = "/root/reference/test"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_literals.json")

NUM = r"[-+]?(?:\d+\.?\d*(?:[eE][-+]?\d+)?|\.\d+)"


def literals(path):
    """Yield (line_number, name, scale, rows) for every `name = [scale *] [ numeric rows ]` block."""
    lines = open(path, encoding="utf-8").read().split("\n")
    i = 0
    while i < len(lines):
        m = re.match(r"^\s*(\w+)\s*=\s*(?:(" + NUM + r")\s*\*\s*)?\[\s*$", lines[i])
        if not m:
            i += 1
            continue
        name, scale = m.group(1), float(m.group(2)) if m.group(2) else 1.0
        rows, j, ok = [], i + 1, True
        while j < len(lines) and not lines[j].strip().startswith("]"):
            txt = lines[j].strip().rstrip(";")
            if txt and not txt.startswith("#"):
                toks = txt.split()
                if not all(re.fullmatch(NUM, t) for t in toks):
                    ok = False
                    break
                rows.append([float(t) for t in toks])
            j += 1
        if ok and rows and len({len(r) for r in rows}) == 1:
            yield i + 1, name, scale, rows
        i = j + 1


def main():
    out = {}
    for fn in ("test_matrixlu.jl", "test_matrixluci.jl"):
        for line, name, scale, rows in literals(os.path.join(REF, fn)):
            key = f"{fn}:{line}:{name}"
            out[key] = {"scale": scale, "rows": rows}
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    for k, v in out.items():
        print(k, len(v["rows"]), "x", len(v["rows"][0]), "scale", v["scale"])


if __name__ == "__main__":
    main()
