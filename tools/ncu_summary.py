#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page) into the handful of metrics DESIGN.md / profiles/ quote."""
import csv
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_warps',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'lts__t_bytes.sum', 'smsp__cycles_active.avg',
        'sm__cycles_elapsed.max', 'sm__cycles_active.avg']


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("kernel:", r[hdr.index("Kernel Name")][:90])
        for w in WANT:
            if w in hdr:
                print(f"  {w:70s} {r[hdr.index(w)]:>16s} {units[hdr.index(w)]}")
        for i, h in enumerate(hdr):
            if "warp_issue_stalled" in h and h.endswith("_per_warp_active.pct"):
                try:
                    v = float(r[i])
                except ValueError:
                    continue
                if v >= 3.0:
                    print(f"  stall {h.split('stalled_')[1].split('_per_warp')[0]:40s} {v:8.1f} %")


if __name__ == "__main__":
    for p in sys.argv[1:]:
        print("==", p)
        main(p)
