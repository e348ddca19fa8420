"""Measure the rollout bookkeeping kernels on one B200.

Same method as bench.py: CUDA-graph replays over rotating replicas larger than L2 (cold L2), CUDA events on
the launching stream.  Beside each kernel: the reference's literal torch lines (A2CAgent_MA.py:34-56 / the
rl_games GAE loop) on the same device, eager, and the C oracle on the host cores.
Algorithmic bytes: step = 45 B/row with value bootstrap (R rewards 4, dones 8, time_outs 1, values 4,
current_* 12; W mb_rewards 4, dones_out 1, current_* 12) -- the done list is < 1 %; GAE = 17 B/row/step
(R mb_rewards 4, mb_values 4, mb_dones 1; W mb_advs 4, mb_returns 4).

    python tools/bench_rollout.py [--rows 16384] [--horizon 16] [--json out.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ige_loader import load_package  # noqa: E402

STEP_BYTES, GAE_BYTES = 45, 17


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6551.0


def script(rows, A, dev, seed):
    g = torch.Generator().manual_seed(seed)
    d = (torch.rand((rows // A,), generator=g) < 0.02).repeat_interleave(A)
    return dict(rewards=torch.randn((rows,), generator=g).to(dev), values=torch.randn((rows,), generator=g).to(dev),
                dones=d.long().to(dev), time_outs=(d & (torch.rand((rows,), generator=g) < 0.3)).to(torch.uint8).to(dev))


def time_graph(g, iters, per_replay):
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (iters * per_replay)  # us per launch


def capture(fn):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return g


def torch_step(st, sc, A, gamma=0.99, scale=0.01):
    """A2CAgent_MA.py:34-56 literally (torch eager on the device; the len() host sync left out)."""
    rewards = sc["rewards"].unsqueeze(1)
    shaped = torch.clamp((rewards + 0.0) * scale, -np.inf, np.inf)
    shaped += gamma * sc["values"].unsqueeze(1) * sc["time_outs"].unsqueeze(1).float()
    st["mb_rewards"][...] = shaped
    st["cr"] += rewards
    st["cs"] += shaped
    st["cl"] += 1
    all_done = sc["dones"].nonzero(as_tuple=False)
    env_done = all_done[::A]
    st["means"] = [torch.mean(x[env_done].float(), dim=0) for x in (st["cr"], st["cs"], st["cl"])]
    nd = 1.0 - sc["dones"].float()
    st["cr"] = st["cr"] * nd.unsqueeze(1)
    st["cs"] = st["cs"] * nd.unsqueeze(1)
    st["cl"] = st["cl"] * nd


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, nargs="*", default=[16384, 262144, 4194304])
    ap.add_argument("--agents", type=int, default=2)
    ap.add_argument("--horizon", type=int, default=16)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    pkg = load_package()
    dev = torch.device("cuda:0")
    pk = peak()
    out = []
    for rows in args.rows:
        A, Hz = args.agents, args.horizon
        # replicas: each bookkeeper owns (H rows of mb_*, current_*); rotate enough of them to exceed the 126 MiB L2
        per = rows * (STEP_BYTES + 8 * 0) * 1.0
        reps = int(min(64, max(2, (160 << 20) // max(1, rows * STEP_BYTES * Hz) + 1)))
        bks = [pkg.rollout.RolloutBookkeeperMA(rows, A, Hz, dev, scale_value=0.01) for _ in range(reps)]
        scs = [[script(rows, A, dev, 7 * r + n) for n in range(Hz)] for r in range(reps)]

        def steps():
            for bk, sc in zip(bks, scs):
                bk.begin_rollout()
                for n in range(Hz):
                    bk.after_env_step(sc[n]["rewards"], sc[n]["dones"], time_outs=sc[n]["time_outs"], values=sc[n]["values"])

        g = capture(steps)
        us_step = time_graph(g, 20, reps * Hz)
        vals = [torch.stack([s["values"] for s in sc]).contiguous() for sc in scs]
        lasts = [torch.randn((rows,), device=dev) for _ in range(reps)]

        def gae():
            for bk, v, l in zip(bks, vals, lasts):
                bk.discount_values(l, v)

        g2 = capture(gae)
        us_gae = time_graph(g2, 20, reps)
        # torch eager, same device
        st = dict(mb_rewards=torch.zeros((rows, 1), device=dev), cr=torch.zeros((rows, 1), device=dev),
                  cs=torch.zeros((rows, 1), device=dev), cl=torch.zeros((rows,), device=dev))
        for _ in range(5):
            torch_step(st, scs[0][0], A)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        k = 50
        for i in range(k):
            torch_step(st, scs[i % reps][i % Hz], A)
        torch.cuda.synchronize()
        us_torch = (time.perf_counter() - t0) * 1e6 / k
        rec = {"rows": rows, "agents": A, "horizon": Hz, "replicas": reps,
               "rollout_step_us": us_step, "rollout_step_gbs": rows * STEP_BYTES / us_step * 1e-3,
               "rollout_step_frac": rows * STEP_BYTES / us_step * 1e-3 / pk,
               "rows_per_s": rows / us_step * 1e6,
               "gae_us": us_gae, "gae_gbs": rows * Hz * GAE_BYTES / us_gae * 1e-3,
               "gae_frac": rows * Hz * GAE_BYTES / us_gae * 1e-3 / pk,
               "torch_eager_same_device_step_us": us_torch, "peak_gbs": pk}
        print(json.dumps(rec))
        out.append(rec)
        del bks, scs, g, g2
        torch.cuda.empty_cache()
    if args.json:
        json.dump(out, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
