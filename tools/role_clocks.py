#!/usr/bin/env python
"""Per-role cycle breakdown of the pipelined locomotion kernel.  Needs the debug build:
    python isaacgymenvs-ma_b200/build.py -DIGE_ROLE_CLOCKS
    python tools/role_clocks.py --task Ant --num-envs 1048576
(rebuild without the flag afterwards: the counters cost a few percent)."""
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
    ap.add_argument("--steps", type=int, default=8)
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
    for s in range(a.replicas):
        envs[s].post_physics_step()
    buf = (ctypes.c_ulonglong * 64)()
    lib.ige_debug_role_clocks(None, 1)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for s in range(a.steps):
        envs[s % a.replicas].post_physics_step()
    ev1.record()
    torch.cuda.synchronize()
    lib.ige_debug_role_clocks(buf, 0)
    v = [[buf[r * 8 + c] for c in range(8)] for r in range(8)]
    print(f"{a.task} {a.num_envs}: {ev0.elapsed_time(ev1) * 1e3 / a.steps:.1f} us/step (eager, instrumented)")
    names = ["HEADING", "VELOCITY", "ANGLES", "REWARD", "ELEM*"]
    for r, nm in enumerate(names):
        tiles = max(v[r][7], 1)
        if nm == "REWARD":
            print(f"  {nm:9s} per tile: work {v[r][1] / tiles:8.0f}  barrier {v[r][2] / tiles:8.0f} | per CTA-warp: "
                  f"prologue {v[r][0] * a.steps / tiles:8.0f}/tiles-per-cta  tail(list) {v[r][3]:d} total, lifetime {v[r][4]:d} total, tiles {tiles}")
        else:
            print(f"  {nm:9s} per tile: load-wait {v[r][0] / tiles:8.0f}  work {v[r][1] / tiles:8.0f}  barrier {v[r][2] / tiles:8.0f}  "
                  f"issue {v[r][4] / tiles:8.0f}  loop-total {v[r][5] / tiles:8.0f}  tiles {tiles}")


    # appended by the arrival-time instrumentation
    return v, names


def report_arrivals(v, names):
    tiles = max(v[3][7], 1)
    print("  last to arrive at the tile barrier (share of tiles) / mean wait for the last arrival (cycles):")
    for r, nm in enumerate(names):
        if nm == "REWARD":
            print(f"    {nm:9s} last on {100.0 * v[5][r] / tiles:5.1f}% of tiles, late by {v[6][r] / tiles:7.0f} cyc/tile on average; "
                  f"idle before the barrier {v[5][7] / tiles:7.0f} cyc/tile")
        else:
            n = 1 if nm != "ELEM*" else max(1, round(v[4][7] / tiles))
            print(f"    {nm:9s} last on {100.0 * v[5][r] / tiles:5.1f}% of tiles, waits {v[6][r] / tiles / n:7.0f} cyc/tile")


if __name__ == "__main__":
    report_arrivals(*main())
