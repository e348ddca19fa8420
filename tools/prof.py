#!/usr/bin/env python
"""Tiny driver for ncu / compute-sanitizer: a few eager post_physics_step calls of one task on
rotating state replicas (so inputs do not sit in L2).  Usage:
    python tools/prof.py --task Ant --num-envs 16384 --steps 4 --replicas 4
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ige_loader import load_package  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="Ant")
    ap.add_argument("--num-envs", type=int, default=16384)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--replicas", type=int, default=4)
    ap.add_argument("--epilogue", action="store_true")
    ap.add_argument("--hot", action="store_true", help="hot_step() (incl. the pre-physics reset of ShadowHand)")
    ap.add_argument("--fall-frac", type=float, default=0.02)
    a = ap.parse_args()
    pkg = load_package()
    envs = []
    for r in range(a.replicas):
        st = pkg.synthetic.make_state(pkg.factory._SYN_NAME[a.task], a.num_envs, seed=r, reset_prob=0.02, **({"fall_frac": a.fall_frac} if a.task in ("Ant", "Humanoid") else {}))
        e = pkg.factory.from_state(a.task, st, "cuda:0", rand_mode="philox", seed=1, fuse_epilogue=a.epilogue)
        e.sim.record = False
        envs.append(e)
    torch.cuda.synchronize()
    for s in range(a.steps):
        (envs[s % a.replicas].hot_step if a.hot else envs[s % a.replicas].post_physics_step)()
    torch.cuda.synchronize()
    print("ok", a.task, a.num_envs, a.steps)


if __name__ == "__main__":
    main()
