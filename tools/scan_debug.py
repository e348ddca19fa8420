#!/usr/bin/env python
"""Timeline of the reset-list scanner CTA (ige::list_scanner_cta).  Needs the debug build:
    python isaacgymenvs-ma_b200/build.py -DIGE_SCAN_DEBUG
    python tools/scan_debug.py --task Ant --num-envs 1048576
Times are %globaltimer nanoseconds relative to the first worker CTA's start."""
import argparse
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ige_loader import load_package  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="Ant")
    ap.add_argument("--num-envs", type=int, default=1048576)
    ap.add_argument("--replicas", type=int, default=4)
    a = ap.parse_args()
    pkg = load_package()
    lib = pkg._lib.lib()
    envs = []
    for r in range(a.replicas):
        st = pkg.synthetic.make_state(a.task.lower(), a.num_envs, seed=r, reset_prob=0.02, fall_frac=0.02)
        e = pkg.factory.from_state(a.task, st, "cuda:0", rand_mode="philox", seed=1)
        e.sim.record = False
        envs.append(e)
    for rep in range(3):
        for e in envs:
            e.post_physics_step()
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * 640)()
    for trial in range(2):
        lib.ige_debug_scan(None, 1)
        envs[trial].post_physics_step()
        torch.cuda.synchronize()
        lib.ige_debug_scan(buf, 0)
        v = [[buf[r * 8 + c] for c in range(8)] for r in range(80)]
        t0 = v[0][0]
        print(f"{a.task} {a.num_envs} trial {trial}: workers end {v[0][1] - t0} ns, scanner start {v[0][2] - t0} ns")
        rows = [(u, v[1 + u]) for u in range(79) if v[1 + u][0]]
        for u, r in rows[:12] + rows[-6:]:
            print(f"  unit {u:3d}: poll start {r[0] - t0:8d}  polls {r[1]:5d}  valid {r[2] - t0:8d}  scanned {r[5] - t0:8d}  spun {r[6] - t0:8d}  chain {r[3] - t0:8d}  ids {r[4] - t0:8d}")


if __name__ == "__main__":
    main()
