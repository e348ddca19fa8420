"""Golden fixtures for the MA Franka OSC torque pre-physics step.

Calls the UNMODIFIED reference ``Franka*MA.pre_physics_step`` / ``_compute_osc_torques``
(tasks/franka_reach_MA.py:770-818, tasks/franka_PPMA.py:897-956) on objects built by tests/ref_harness.py, with
``_get_mm`` / ``_get_j_eef`` (the PhysX mass-matrix / Jacobian tensors, franka_PPMA.py:1041-1061) replaced by
synthetic well-conditioned tensors: mass matrices SPD, Jacobians full row rank.

Container-only (needs /root/reference):  python tests/golden/gen_golden_osc.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as H  # noqa: E402
import ref_harness  # noqa: E402

EFFORT_LIMITS = [87.0, 87.0, 87.0, 87.0, 12.0, 12.0, 12.0]
REF_CMD_LIMIT = {"reach": [0.1, 0.1, 0.1, 0.5, 0.5, 0.5], "collect": [0.6] * 6, "combine": [0.55] * 6, "pp": [0.55] * 6}


def synthetic_dynamics(rows, seed):
    g = torch.Generator().manual_seed(seed)
    G = torch.randn((rows, 7, 7), generator=g)
    d = torch.rand((rows, 7), generator=g) * 1.5 + 0.5
    mm = torch.diag_embed(d) + 0.1 * G @ G.transpose(1, 2)          # SPD, condition number O(10)
    j = 0.2 * torch.randn((rows, 6, 7), generator=g)
    j[:, :, :6] += torch.eye(6)                                      # full row rank
    return mm.contiguous(), j.contiguous()


def generate(variant, n, A, seed):
    syn = H.pkg.synthetic
    cfg = syn.franka_ma_cfg(variant, A, 2)
    st = syn.make_state("franka_" + variant, n, seed=seed, num_agents=A, num_targets=2)
    obj = ref_harness.build_franka_ma(st, cfg)
    del obj.pre_physics_step                                          # back to the reference's own method
    rows = n * A
    mm, j = synthetic_dynamics(rows, seed + 100)
    obj._get_mm, obj._get_j_eef = (lambda: mm), (lambda: j)
    obj.control_type = "osc"
    obj.action_scale = 1.0                                            # cfg/task/FrankaPPMA.yaml actionScale
    # gentle motion so that about half of the torques stay inside the effort limits (the clamp is tested either way)
    obj._rigid_body_state[:, :, 7:] *= 0.05
    obj._dof_state[..., 1] *= 0.05
    obj.kp = torch.tensor([150.0] * 6)
    obj.kd = 2 * torch.sqrt(obj.kp)
    obj.kp_null = torch.tensor([10.0] * 7)
    obj.kd_null = 2 * torch.sqrt(obj.kp_null)
    # the reference's own per-variant command limits (franka_reach_MA.py:219 is NOT uniform over the six axes;
    # franka_collect_MA.py:228, franka_combine_MA.py / franka_PPMA.py:234)
    obj.cmd_limit = torch.tensor(REF_CMD_LIMIT[variant]).unsqueeze(0)
    obj._franka_effort_limits = torch.tensor(EFFORT_LIMITS + [70.0, 70.0])
    obj._arm_control = obj._effort_control.view(-1, 9)[:, :7]
    obj._gripper_control = obj._pos_control.view(-1, 9)[:, -2:]
    obj.states["eef_vel"] = obj._eef_state[:, :, 7:]
    g = torch.Generator().manual_seed(seed + 7)
    raw = torch.rand((rows, cfg["num_actions"]), generator=g) * 3.0 - 1.5
    raw[1::2, :6] *= 0.08      # every other row: small arm commands, so that a good share of the torques is NOT clamped
    actions = torch.clamp(raw, -1.0, 1.0)                             # VecTask.step, vec_task.py:373
    blob = {"in_actions_raw": raw.numpy().copy(), "in_mm": mm.numpy().copy(), "in_j_eef": j.numpy().copy(),
            "in_dof_state": obj._dof_state.numpy().copy(), "in_rigid_body_state": obj._rigid_body_state.numpy().copy(),
            "in_pos_control": obj._pos_control.numpy().copy(), "in_effort_control": obj._effort_control.numpy().copy()}
    type(obj).pre_physics_step(obj, actions)                          # <- the reference's own code
    blob.update({
        "out_actions": obj.actions.numpy().copy(), "out_effort_control": obj._effort_control.numpy().copy(),
        "out_pos_control": obj._pos_control.numpy().copy(),
        "cfg_variant": variant, "cfg_num_agents": A, "cfg_num_envs": n, "cfg_num_actions": cfg["num_actions"],
        "cfg_eef_body_start": cfg["eef_body_start"], "cfg_eef_body_step": cfg["eef_body_step"],
        "cfg_num_bodies": obj._rigid_body_state.shape[1],
        "cfg_kp": obj.kp.numpy(), "cfg_kd": obj.kd.numpy(), "cfg_kp_null": obj.kp_null.numpy(),
        "cfg_kd_null": obj.kd_null.numpy(), "cfg_cmd_limit": obj.cmd_limit.flatten().numpy(),
        "cfg_action_scale": obj.action_scale, "cfg_effort_limits": np.array(EFFORT_LIMITS, np.float32),
        "cfg_default_dof_pos": obj.franka_default_dof_pos.numpy()[:7],
        "cfg_gripper_lower": obj.franka_dof_lower_limits.numpy()[-2:],
        "cfg_gripper_upper": obj.franka_dof_upper_limits.numpy()[-2:],
    })
    return blob


FIXTURES = [("pp", 48, 2, 0), ("reach", 40, 2, 1), ("collect", 21, 3, 2), ("combine", 33, 2, 3)]


def main():
    for variant, n, A, seed in FIXTURES:
        blob = generate(variant, n, A, seed)
        path = os.path.join(HERE, f"osc_{variant}_n{n}x{A}_s{seed}.npz")
        np.savez_compressed(path, **blob)
        eff = blob["out_effort_control"]
        print(f"{os.path.basename(path)}: {os.path.getsize(path)} B, |u| max {np.abs(eff).max():.2f}, "
              f"clamped {float((np.abs(eff.reshape(-1, 9)[:, :6]) >= np.array(EFFORT_LIMITS[:6]) - 1e-6).mean()):.2f}")


if __name__ == "__main__":
    main()
