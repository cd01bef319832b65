#!/usr/bin/env python3
"""Aggregate the warp-stall samples of an `ncu --set full --import-source on` report per CUDA source line.

    python tools/ncu_hotlines.py gpurun_out/prof_c3.ncu-rep [top_n]
"""
import csv
import io
import subprocess
import sys
from collections import defaultdict


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
    hdr = rows[hdr_i]
    col = {n: i for i, n in enumerate(hdr)}
    # the first "Source" column is the CUDA line, the second the SASS instruction
    src_i = hdr.index("Source")
    sass_i = hdr.index("Source", src_i + 1)
    stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    per_line = defaultdict(lambda: defaultdict(float))
    text = {}
    total = 0.0
    for r in rows[hdr_i + 1:]:
        if len(r) != len(hdr):
            continue
        try:
            n = float(r[col["# Samples"]] or 0)
        except ValueError:
            continue
        key = r[0]
        text[key] = r[src_i].strip()[:110]
        per_line[key]["samples"] += n
        per_line[key]["inst"] += float(r[col["Instructions Executed"]] or 0)
        total += n
        for s in stall_cols:
            try:
                per_line[key][s] += float(r[col[s]] or 0)
            except ValueError:
                pass
    print(f"# total samples {total:.0f}")
    for key, d in sorted(per_line.items(), key=lambda kv: -kv[1]["samples"])[:top]:
        stalls = sorted(((d[s], s[6:]) for s in stall_cols if d[s] > 0), reverse=True)[:3]
        st = " ".join(f"{n}:{v / max(d['samples'], 1):.0%}" for v, n in stalls)
        print(f"{d['samples'] / total:6.1%} L{key:>5s} inst {d['inst']:>10.0f}  [{st}]  {text[key]}")


if __name__ == "__main__":
    main()
