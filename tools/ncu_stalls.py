#!/usr/bin/env python
"""Per-source-line stall samples of one kernel from an .ncu-rep (needs -lineinfo + --import-source on):
    python tools/ncu_stalls.py gpurun_out/x.ncu-rep [--sass] [--top 40]
Aggregates the warp-stall sampling columns of `ncu --page source --csv` by CUDA source line (or prints the hottest
SASS instructions) and names the dominant stall reason of each."""
import csv
import subprocess
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    view = "sass" if "--sass" in sys.argv else "cuda,sass"
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", view],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    si = hdr.index("# Samples")
    src_i = hdr.index("Source")
    total = 0
    agg = defaultdict(lambda: [0, defaultdict(int), ""])
    cur_line = "?"
    for r in rows[hdr_i + 1:]:
        if len(r) <= si:
            continue
        try:
            n = int(r[si])
        except ValueError:
            continue
        key = r[src_i].strip()[:110]
        a = agg[key]
        a[0] += n
        total += n
        for c in stall_cols:
            try:
                a[1][hdr[c]] += int(r[c])
            except ValueError:
                pass
    print(f"total samples {total}")
    for key, (n, st, _) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        if n == 0:
            break
        reasons = ", ".join(f"{k[6:]} {v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3] if v)
        print(f"{n:6d} {100.0 * n / max(total, 1):5.1f}%  {key:110s} | {reasons}")


if __name__ == "__main__":
    main()
