"""Drive the MA Franka OSC pre-physics step through the CPU oracle (host pointers) or the
CUDA library (device pointers) from one IgeFrankaOscArgs struct; inputs / outputs as in tests/golden/gen_golden_osc.py."""
import glob
import os

import numpy as np
import torch

import harness as H


def golden_files():
    return sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "osc_*.npz")))


def fill_args(blob, ptr):
    """ptr(name) -> address of the (host or device) buffer called `name`."""
    abi = H.pkg._abi
    a = abi.IgeFrankaOscArgs()
    a.num_envs, a.num_agents = int(blob["cfg_num_envs"]), int(blob["cfg_num_agents"])
    a.num_actions = int(blob["cfg_num_actions"])
    a.has_gripper = int(str(blob["cfg_variant"]) != "reach")
    a.num_bodies = int(blob["cfg_num_bodies"])
    a.eef_body_start, a.eef_body_step = int(blob["cfg_eef_body_start"]), int(blob["cfg_eef_body_step"])
    a.clip_actions, a.action_scale = 1.0, float(blob["cfg_action_scale"])
    for name in ("cmd_limit", "kp", "kd", "kp_null", "kd_null", "default_dof_pos", "effort_limits", "gripper_lower",
                 "gripper_upper"):
        abi.fill_array(getattr(a, name), np.asarray(blob["cfg_" + name], np.float32).tolist())
    for field, name in (("actions_in", "in_actions_raw"), ("actions", "actions"), ("mm", "in_mm"), ("j_eef", "in_j_eef"),
                        ("dof_state", "in_dof_state"), ("rigid_body_state", "in_rigid_body_state"),
                        ("effort_control", "in_effort_control"), ("pos_control", "in_pos_control")):
        setattr(a, field, ptr(name))
    return a


def oracle_osc(blob):
    bufs = {k: np.ascontiguousarray(v, np.float32).copy() for k, v in blob.items() if k.startswith("in_")}
    bufs["actions"] = np.zeros_like(bufs["in_actions_raw"])
    H.oracle.call("franka_ma_pre_physics", fill_args(blob, lambda k: bufs[k].ctypes.data))
    return {"out_actions": bufs["actions"], "out_effort_control": bufs["in_effort_control"],
            "out_pos_control": bufs["in_pos_control"]}


def gpu_osc(blob, device="cuda:0"):
    bufs = {k: torch.from_numpy(np.ascontiguousarray(v, np.float32)).to(device) for k, v in blob.items()
            if k.startswith("in_")}
    bufs["actions"] = torch.zeros_like(bufs["in_actions_raw"])
    H.pkg._lib.call("ige_franka_ma_pre_physics", fill_args(blob, lambda k: bufs[k].data_ptr()),
                    stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return {"out_actions": bufs["actions"].cpu().numpy(), "out_effort_control": bufs["in_effort_control"].cpu().numpy(),
            "out_pos_control": bufs["in_pos_control"].cpu().numpy()}


def torque_error(got, ref):
    """max |d| / max(1, |ref|) over the arm torques."""
    g, r = got["out_effort_control"].astype(np.float64), ref["out_effort_control"].astype(np.float64)
    return float((np.abs(g - r) / np.maximum(1.0, np.abs(r))).max())


def rows_over(got, ref, bar=1e-5):
    """(rows whose worst arm torque is further than `bar` * max(1, |ref|) from ref, rows): how much of the fixture
    actually needs the relaxed (1e-4) bar of the two un-pivoted inverses."""
    g, r = got["out_effort_control"].astype(np.float64), ref["out_effort_control"].astype(np.float64)
    err = (np.abs(g - r) / np.maximum(1.0, np.abs(r))).reshape(-1, g.shape[-1]).max(axis=1)
    return int((err > bar).sum()), int(err.size)


def float64_truth(blob):
    """_compute_osc_torques (franka_PPMA.py:897-929) evaluated in float64 on the same inputs: the yardstick for the
    fp32 rounding of BOTH the reference (LAPACK LU) and this repo (Gauss-Jordan)."""
    n, A = int(blob["cfg_num_envs"]), int(blob["cfg_num_agents"])
    rows = n * A
    f = lambda k: np.asarray(blob[k], np.float64)
    act = np.clip(f("in_actions_raw"), -1.0, 1.0)
    mm, J = f("in_mm"), f("in_j_eef")
    ds = f("in_dof_state").reshape(rows, 9, 2)
    q, qd = ds[:, :7, 0], ds[:, :7, 1]
    rb = f("in_rigid_body_state")
    s, st = int(blob["cfg_eef_body_start"]), int(blob["cfg_eef_body_step"])
    ev = rb[:, s:s + st * A:st, 7:].reshape(rows, 6)
    dpose = act[:, :6] * f("cfg_cmd_limit") / float(blob["cfg_action_scale"])
    mm_inv = np.linalg.inv(mm)
    m_eef = np.linalg.inv(J @ mm_inv @ J.transpose(0, 2, 1))
    u = J.transpose(0, 2, 1) @ m_eef @ (f("cfg_kp") * dpose - f("cfg_kd") * ev)[..., None]
    j_eef_inv = m_eef @ J @ mm_inv
    un = f("cfg_kd_null") * -qd + f("cfg_kp_null") * ((f("cfg_default_dof_pos") - q + np.pi) % (2 * np.pi) - np.pi)
    un = mm @ un[..., None]
    u = u + (np.eye(7) - J.transpose(0, 2, 1) @ j_eef_inv) @ un
    lim = f("cfg_effort_limits")
    u = np.clip(u[..., 0], -lim, lim)
    eff = f("in_effort_control").reshape(rows, 9).copy()
    eff[:, :6] = u[:, :6]
    return {"out_effort_control": eff.reshape(n, 9 * A)}


def torch_fp32(blob):
    """The reference's lines (franka_PPMA.py:897-944) typed out in torch fp32 on the CPU, for sizes without a golden."""
    n, A = int(blob["cfg_num_envs"]), int(blob["cfg_num_agents"])
    rows = n * A
    f = lambda k: torch.from_numpy(np.ascontiguousarray(blob[k], np.float32))
    c = lambda k: torch.tensor(np.asarray(blob[k], np.float32))
    act = torch.clamp(f("in_actions_raw"), -1.0, 1.0)
    _mm, _j_eef = f("in_mm"), f("in_j_eef")
    ds = f("in_dof_state").reshape(rows, 9, 2)
    q, qd = ds[:, :7, 0], ds[:, :7, 1]
    s, st = int(blob["cfg_eef_body_start"]), int(blob["cfg_eef_body_step"])
    eef_vel = f("in_rigid_body_state")[:, s:s + st * A:st, 7:].reshape(rows, -1)
    dpose = act[:, :6] * c("cfg_cmd_limit").unsqueeze(0) / float(blob["cfg_action_scale"])
    mm_inv = torch.inverse(_mm)
    m_eef_inv = _j_eef @ mm_inv @ torch.transpose(_j_eef, 1, 2)
    m_eef = torch.inverse(m_eef_inv)
    u = torch.transpose(_j_eef, 1, 2) @ m_eef @ (c("cfg_kp") * dpose - c("cfg_kd") * eef_vel).unsqueeze(-1)
    j_eef_inv = m_eef @ _j_eef @ mm_inv
    u_null = c("cfg_kd_null") * -qd + c("cfg_kp_null") * ((c("cfg_default_dof_pos") - q + np.pi) % (2 * np.pi) - np.pi)
    u_null = _mm @ u_null.unsqueeze(-1)
    u += (torch.eye(7).unsqueeze(0) - torch.transpose(_j_eef, 1, 2) @ j_eef_inv) @ u_null
    lim = c("cfg_effort_limits").unsqueeze(0)
    u = torch.max(torch.min(u.squeeze(-1), lim), -lim)
    eff = f("in_effort_control").reshape(rows, 9).clone()
    eff[:, :6] = u[:, :6]
    return {"out_effort_control": eff.reshape(n, 9 * A).numpy()}
