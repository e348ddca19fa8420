"""Measure ige_franka_ma_pre_physics on one B200: CUDA-graph replays over rotating input replicas
larger than L2, CUDA events; beside it the reference's torch lines (eager, same device).
Algorithmic bytes: 548 B / agent row (R mm 196, j_eef 168, dof_state 72, eef_vel 24, actions 28; W actions 28,
arm torques 24, gripper targets 8).

    python tools/bench_osc.py [--rows 16384 262144 2097152] [--json out.json]
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

ROW_BYTES = 548


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, nargs="*", default=[16384, 262144, 2097152])
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    pkg = load_package()
    abi, lib = pkg._abi, pkg._lib
    dev = torch.device("cuda:0")
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6551.0
    out = []
    for rows in args.rows:
        A, n, nb = 2, rows // 2, 24
        reps = int(min(48, max(2, (200 << 20) // (rows * ROW_BYTES) + 1)))
        g = torch.Generator(device=dev).manual_seed(0)
        sets = []
        for _ in range(reps):
            G = torch.randn((rows, 7, 7), device=dev, generator=g)
            mm = torch.diag_embed(torch.rand((rows, 7), device=dev, generator=g) * 1.5 + 0.5) + 0.1 * G @ G.transpose(1, 2)
            j = 0.2 * torch.randn((rows, 6, 7), device=dev, generator=g)
            j[:, :, :6] += torch.eye(6, device=dev)
            sets.append(dict(mm=mm.contiguous(), j=j.contiguous(),
                             ds=torch.randn((n, 9 * A, 2), device=dev, generator=g),
                             rb=torch.randn((n, nb, 13), device=dev, generator=g) * 0.05,
                             raw=torch.rand((rows, 7), device=dev, generator=g) * 3 - 1.5,
                             act=torch.zeros((rows, 7), device=dev), eff=torch.zeros((n, 9 * A), device=dev),
                             pos=torch.zeros((n, 9 * A), device=dev)))
            del G
        structs = []
        for s in sets:
            a = abi.IgeFrankaOscArgs()
            a.num_envs, a.num_agents, a.num_actions, a.has_gripper, a.num_bodies = n, A, 7, 1, nb
            a.eef_body_start, a.eef_body_step, a.clip_actions, a.action_scale = 9, 11, 1.0, 1.0
            for name, vals in (("cmd_limit", [0.1] * 6), ("kp", [150.0] * 6), ("kd", [24.494898] * 6),
                               ("kp_null", [10.0] * 7), ("kd_null", [6.3245554] * 7), ("default_dof_pos", [0, 0.2, 0, -2.3, 0, 2.7, 0.8]),
                               ("effort_limits", [87, 87, 87, 87, 12, 12, 12]), ("gripper_lower", [0, 0]), ("gripper_upper", [0.04, 0.04])):
                abi.fill_array(getattr(a, name), vals)
            a.actions_in, a.actions, a.mm, a.j_eef = s["raw"].data_ptr(), s["act"].data_ptr(), s["mm"].data_ptr(), s["j"].data_ptr()
            a.dof_state, a.rigid_body_state = s["ds"].data_ptr(), s["rb"].data_ptr()
            a.effort_control, a.pos_control = s["eff"].data_ptr(), s["pos"].data_ptr()
            structs.append(a)

        def launch_all():
            st = torch.cuda.current_stream().cuda_stream
            for a in structs:
                lib.call("ige_franka_ma_pre_physics", a, stream=st)

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            launch_all()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            launch_all()
        for _ in range(3):
            gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 20
        e0.record()
        for _ in range(iters):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (iters * reps)

        def torch_ref(s):
            mm, J = s["mm"], s["j"]
            act = torch.clamp(s["raw"], -1, 1)
            q, qd = s["ds"].view(rows, 9, 2)[:, :7, 0], s["ds"].view(rows, 9, 2)[:, :7, 1]
            ev = s["rb"][:, 9:9 + 22:11, 7:].reshape(rows, 6)
            mm_inv = torch.inverse(mm)
            m_eef = torch.inverse(J @ mm_inv @ J.transpose(1, 2))
            u = J.transpose(1, 2) @ m_eef @ (150.0 * act[:, :6] * 0.1 - 24.49 * ev).unsqueeze(-1)
            un = 6.32 * -qd + 10.0 * ((0.5 - q + np.pi) % (2 * np.pi) - np.pi)
            un = mm @ un.unsqueeze(-1)
            u += (torch.eye(7, device=dev).unsqueeze(0) - J.transpose(1, 2) @ (m_eef @ J @ mm_inv)) @ un
            return torch.clamp(u.squeeze(-1), -87, 87)

        for _ in range(3):
            torch_ref(sets[0])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        k = 10
        for i in range(k):
            torch_ref(sets[i % reps])
        torch.cuda.synchronize()
        us_torch = (time.perf_counter() - t0) * 1e6 / k
        rec = {"rows": rows, "replicas": reps, "osc_us": us, "osc_gbs": rows * ROW_BYTES / us * 1e-3,
               "osc_frac": rows * ROW_BYTES / us * 1e-3 / peak, "agent_rows_per_s": rows / us * 1e6,
               "torch_eager_same_device_us": us_torch, "peak_gbs": peak}
        print(json.dumps(rec))
        out.append(rec)
        del sets, structs, gr
        torch.cuda.empty_cache()
    if args.json:
        json.dump(out, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
