#!/usr/bin/env python
"""Per-source-line stall samples / executed instructions of an .ncu-rep (source page, cuda+sass view).
Usage: python tools/ncu_lines.py REPORT [top_n] [launch_index]"""
import csv
import subprocess
import sys


def main(path, top=40):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    fname, hdr, lines, seen_kernel = None, None, {}, 0
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            hdr = r
            si, ii = hdr.index("# Samples"), hdr.index("Instructions Executed")
            stall0 = hdr.index("stall_barrier")
            continue
        if hdr is None or not r[0].isdigit():
            continue
        key = (fname, int(r[0]))
        s = int(r[si]) if r[si].isdigit() else 0
        i = int(r[ii]) if r[ii].isdigit() else 0
        st = {}
        for k in range(stall0, min(stall0 + 17, len(r))):
            if r[k].isdigit() and int(r[k]) > 0:
                st[hdr[k].replace("stall_", "")] = int(r[k])
        if key in lines:
            continue  # first launch only
        lines[key] = (s, i, r[1].strip()[:90], st)
    tot_s = sum(v[0] for v in lines.values()) or 1
    tot_i = sum(v[1] for v in lines.values()) or 1
    print(f"total samples {tot_s}, instructions {tot_i}")
    for (f, ln), (s, i, src, st) in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
        top_st = ",".join(f"{k}:{v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{s:6d} {100*s/tot_s:5.1f}%  inst {i:9d} {100*i/tot_i:5.1f}%  {f}:{ln}: {src}   [{top_st}]")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
