"""Synthetic Anymal state (shared by tests/golden/gen_golden_anymal.py and the GPU parity tests; no reference needed)."""
import numpy as np

DT = 0.02
CFG = dict(max_episode_length=int(50 / DT + 0.5), clip_obs=5.0, clip_actions=1.0,
           lin_vel_scale=2.0, ang_vel_scale=0.25, dof_pos_scale=1.0, dof_vel_scale=0.05, action_scale=0.5,
           rew_lin_vel_xy=1.0, rew_ang_vel_z=0.5, rew_torque=-0.000025,
           command_x_range=[-2.0, 2.0], command_y_range=[-1.0, 1.0], command_yaw_range=[-1.0, 1.0],
           num_bodies=17, base_index=0, knee_indices=[2, 6, 10, 14],
           # cfg/task/Anymal.yaml defaultJointAngles in DOF order (LF, LH, RF, RH: HAA, HFE, KFE)
           default_dof_pos=[0.03, 0.4, -0.8, 0.03, -0.4, 0.8, -0.03, 0.4, -0.8, -0.03, -0.4, 0.8])


def make_state(n, seed, reset_prob):
    g = np.random.Generator(np.random.PCG64(seed))
    f = lambda *s: g.standard_normal(s).astype(np.float32)
    root = np.zeros((n, 13), np.float32)
    root[:, :3] = f(n, 3)
    q = f(n, 4)
    half = n // 2
    q[:half] = np.array([0, 0, 0, 1], np.float32) + 0.15 * f(half, 4)       # mostly upright
    root[:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
    root[:, 7:13] = f(n, 6)
    contact = 0.4 * f(n, CFG["num_bodies"], 3)                              # norms straddle the > 1 threshold
    progress = g.integers(0, CFG["max_episode_length"] + 2, n).astype(np.int64)
    progress[: max(1, n // 16)] = CFG["max_episode_length"] - 2             # time-outs this step
    return {
        "root_states": root, "dof_state": f(n, 12, 2) * np.array([0.5, 2.0], np.float32),
        "contact_forces": contact.astype(np.float32), "torques": 20.0 * f(n, 12),
        "actions": g.uniform(-1.5, 1.5, (n, 12)).astype(np.float32),
        "commands": g.uniform(-1, 1, (n, 3)).astype(np.float32),
        "reset_buf": (g.random(n) < reset_prob).astype(np.int64), "progress_buf": progress,
    }


def advance(st, g):
    for k in ("root_states", "dof_state", "contact_forces", "torques"):
        st[k] = (st[k] + 0.1 * g.standard_normal(st[k].shape)).astype(np.float32)
    q = st["root_states"][:, 3:7]
    st["root_states"][:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
    st["actions"] = g.uniform(-1.5, 1.5, st["actions"].shape).astype(np.float32)
