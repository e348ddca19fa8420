"""Seeded synthetic simulator state for the post-physics hot path.

There is no PhysX here (the Isaac Gym binary is absent), so the tensors ``gymtorch.wrap_tensor``
would expose after ``gym.simulate``/``refresh_*`` are *generated*: physically plausible, and
deliberately straddling every branch of the reference code (termination height, joint limits,
episode-length time-outs, a controllable fraction of ``reset_buf == 1``).

Everything is numpy (PCG64, bit-stable across machines) in the gymtorch layouts:
root / rigid-body rows are 13 floats (pos3, quat xyzw4, linvel3, angvel3), DOF rows are
(pos, vel), force sensors 6 floats, contact forces 3 floats
(reference: tasks/ant.py:78-95, tasks/shadow_hand.py:150-175, tasks/franka_reach_MA.py:453-460).
"""
from __future__ import annotations

import math

import numpy as np

F32 = np.float32
I64 = np.int64

# ---------------------------------------------------------------------------------------------
# task constants (reference cfg values)
# ---------------------------------------------------------------------------------------------
_DEG = math.pi / 180.0

CARTPOLE_CFG = dict(reset_dist=3.0, max_episode_length=500, clip_obs=5.0, dt=0.0166)

# assets/mjcf/nv_ant.xml:48-75 (hip +-40 deg, ankles [30,100] / [-100,-30])
ANT_DOF_LOWER = np.array([-40, 30, -40, -100, -40, -100, -40, 30], dtype=np.float64) * _DEG
ANT_DOF_UPPER = np.array([40, 100, 40, -30, 40, -30, 40, 100], dtype=np.float64) * _DEG
ANT_CFG = dict(
    max_episode_length=1000, dof_vel_scale=0.2, contact_force_scale=0.1, power_scale=1.0,
    heading_weight=0.5, up_weight=0.1, actions_cost_scale=0.005, energy_cost_scale=0.05,
    joints_at_limit_cost_scale=0.1, death_cost=-2.0, termination_height=0.31, dt=0.0166,
    clip_obs=float("inf"), start_height=0.44, num_dof=8, num_sensors=4, num_obs=60)

# assets/mjcf/nv_humanoid.xml joint ranges in DOF order (21) and motor gears :141-161
HUMANOID_DOF_RANGE_DEG = [
    (-45, 45), (-75, 30), (-35, 35),                                  # abdomen z,y,x
    (-45, 15), (-60, 35), (-120, 45), (-160, 2), (-50, 50), (-50, 50),  # right leg
    (-45, 15), (-60, 35), (-120, 45), (-160, 2), (-50, 50), (-50, 50),  # left leg
    (-90, 70), (-90, 70), (-90, 50),                                  # right arm
    (-90, 70), (-90, 70), (-90, 50)]                                  # left arm
HUMANOID_DOF_LOWER = np.array([r[0] for r in HUMANOID_DOF_RANGE_DEG], dtype=np.float64) * _DEG
HUMANOID_DOF_UPPER = np.array([r[1] for r in HUMANOID_DOF_RANGE_DEG], dtype=np.float64) * _DEG
HUMANOID_MOTOR_EFFORTS = np.array(
    [67.5, 67.5, 67.5, 45.0, 45.0, 135.0, 90.0, 22.5, 22.5, 45.0, 45.0, 135.0, 90.0, 22.5, 22.5,
     67.5, 67.5, 45.0, 67.5, 67.5, 45.0], dtype=np.float64)
HUMANOID_CFG = dict(
    max_episode_length=1000, dof_vel_scale=0.1, angular_velocity_scale=0.25,
    contact_force_scale=0.01, power_scale=1.0, heading_weight=0.5, up_weight=0.1,
    actions_cost_scale=0.01, energy_cost_scale=0.05, joints_at_limit_cost_scale=0.25,
    death_cost=-1.0, termination_height=0.8, dt=0.0166, clip_obs=float("inf"),
    start_height=1.34, num_dof=21, num_sensors=2, num_obs=108)


def _rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def _unit_quats(rng, n):
    """50 % uniformly random orientations, 50 % small yaw rotations (|angle| <= 0.3 rad)."""
    q = rng.standard_normal((n, 4))
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    small = rng.random(n) < 0.5
    ang = rng.uniform(-0.3, 0.3, n)
    qs = np.stack([np.zeros(n), np.zeros(n), np.sin(ang / 2), np.cos(ang / 2)], -1)
    q[small] = qs[small]
    return q.astype(F32)


def _progress(rng, n, max_len):
    p = rng.integers(0, max_len + 2, n)
    pin = rng.random(n) < 0.01
    p[pin] = max_len - 2          # +1 in the step hits max_len - 1 exactly
    return p.astype(I64)


def _resets(rng, n, prob):
    if prob >= 1.0:
        return np.ones(n, I64)
    if prob <= 0.0:
        return np.zeros(n, I64)
    return (rng.random(n) < prob).astype(I64)


# ---------------------------------------------------------------------------------------------
# Cartpole
# ---------------------------------------------------------------------------------------------
def make_cartpole(num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    rng = _rng(seed)
    n = num_envs
    st = {}
    st["dof_state"] = (rng.standard_normal((n, 2, 2)) * np.array([[3.5, 2.0], [1.2, 2.0]])
                       ).astype(F32).reshape(n * 2, 2)
    st["progress_buf"] = _progress(rng, n, CARTPOLE_CFG["max_episode_length"])
    st["reset_buf"] = _resets(rng, n, reset_prob)
    st["obs_buf"] = np.zeros((n, 4), F32)
    st["rew_buf"] = np.zeros(n, F32)
    return st


def advance_cartpole(st, rng):
    n = st["progress_buf"].shape[0]
    st["dof_state"] = (rng.standard_normal((n, 2, 2)) * np.array([[3.5, 2.0], [1.2, 2.0]])
                       ).astype(F32).reshape(n * 2, 2)


# ---------------------------------------------------------------------------------------------
# Ant / Humanoid (same structure)
# ---------------------------------------------------------------------------------------------
def _heights(rng, n, th, fall_frac):
    """Torso heights straddling the termination height; `fall_frac` pins the fallen fraction."""
    if fall_frac is None:
        return rng.uniform(th - 0.15, th + 0.5, n)
    z = rng.uniform(th + 1e-3, th + 0.5, n)
    fallen = rng.random(n) < fall_frac
    z[fallen] = rng.uniform(th - 0.15, th - 1e-3, int(fallen.sum()))
    return z


def _make_locomotion(cfg, lower, upper, num_envs, seed, reset_prob, with_dof_force, fall_frac=None):
    rng = _rng(seed)
    n, d = num_envs, cfg["num_dof"]
    th, dt = cfg["termination_height"], cfg["dt"]
    st = {}
    root = np.zeros((n, 13), F32)
    root[:, 0:2] = rng.normal(0.0, 3.0, (n, 2))
    root[:, 2] = _heights(rng, n, th, fall_frac)
    root[:, 3:7] = _unit_quats(rng, n)
    root[:, 7:13] = rng.standard_normal((n, 6))
    st["root_states"] = root
    init = np.zeros((n, 13), F32)
    init[:, 0:2] = rng.normal(0.0, 3.0, (n, 2))
    init[:, 2] = cfg["start_height"]
    init[:, 6] = 1.0
    st["initial_root_states"] = init
    lo, hi = lower.astype(F32), upper.astype(F32)
    span = (hi - lo).astype(np.float64)
    dof = np.zeros((n, d, 2), F32)
    dof[:, :, 0] = rng.uniform(lo - 0.05 * span, hi + 0.05 * span, (n, d))
    dof[:, :, 1] = rng.normal(0.0, 2.0, (n, d))
    st["dof_state"] = dof.reshape(n * d, 2)
    st["dof_limits_lower"], st["dof_limits_upper"] = lo, hi
    # ant.py:96-99: zero clamped into the limits, tiled over envs
    init_dof = np.where(lo > 0, lo, np.where(hi < 0, hi, 0.0)).astype(F32)
    st["initial_dof_pos"] = np.tile(init_dof, (n, 1))
    st["vec_sensor_tensor"] = rng.normal(0.0, 5.0, (n, cfg["num_sensors"] * 6)).astype(F32)
    if with_dof_force:
        st["dof_force_tensor"] = rng.normal(0.0, 20.0, (n, d)).astype(F32)
    st["actions"] = rng.uniform(-1.0, 1.0, (n, d)).astype(F32)
    st["progress_buf"] = _progress(rng, n, cfg["max_episode_length"])
    st["reset_buf"] = _resets(rng, n, reset_prob)
    st["randomize_buf"] = np.zeros(n, I64)
    st["targets"] = np.tile(np.array([1000.0, 0.0, 0.0], F32), (n, 1))
    st["inv_start_rot"] = np.tile(np.array([0.0, 0.0, 0.0, 1.0], F32), (n, 1))
    st["basis_vec0"] = np.tile(np.array([1.0, 0.0, 0.0], F32), (n, 1))
    st["basis_vec1"] = np.tile(np.array([0.0, 0.0, 1.0], F32), (n, 1))
    # potentials from the previous synthetic position (pos - vel*dt), fp32 like the reference
    prev = root[:, 0:3] - root[:, 7:10] * F32(dt)
    tt = st["targets"] - prev
    tt[:, 2] = 0.0
    pot = (-np.sqrt((tt.astype(np.float64) ** 2).sum(-1)) / dt).astype(F32)
    st["potentials"] = pot
    st["prev_potentials"] = pot.copy()
    st["up_vec"] = st["basis_vec1"].copy()
    st["heading_vec"] = st["basis_vec0"].copy()
    st["obs_buf"] = np.zeros((n, cfg["num_obs"]), F32)
    st["rew_buf"] = np.zeros(n, F32)
    return st


def _advance_locomotion(cfg, st, rng, with_dof_force):
    """Stand-in for gym.simulate + refresh: integrate the torso, redraw everything else."""
    n, d = st["progress_buf"].shape[0], cfg["num_dof"]
    root = st["root_states"]
    root[:, 0:3] = root[:, 0:3] + root[:, 7:10] * F32(cfg["dt"])
    root[:, 2] = rng.uniform(cfg["termination_height"] - 0.15, cfg["termination_height"] + 0.5, n)
    root[:, 3:7] = _unit_quats(rng, n)
    root[:, 7:13] = rng.standard_normal((n, 6))
    lo, hi = st["dof_limits_lower"], st["dof_limits_upper"]
    span = (hi - lo).astype(np.float64)
    dof = st["dof_state"].reshape(n, d, 2)
    dof[:, :, 0] = rng.uniform(lo - 0.05 * span, hi + 0.05 * span, (n, d))
    dof[:, :, 1] = rng.normal(0.0, 2.0, (n, d))
    st["vec_sensor_tensor"][:] = rng.normal(0.0, 5.0, st["vec_sensor_tensor"].shape)
    if with_dof_force:
        st["dof_force_tensor"][:] = rng.normal(0.0, 20.0, (n, d))
    st["actions"][:] = rng.uniform(-1.0, 1.0, (n, d))


def make_ant(num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    return _make_locomotion(ANT_CFG, ANT_DOF_LOWER, ANT_DOF_UPPER, num_envs, seed, reset_prob, False, fall_frac)


def advance_ant(st, rng):
    _advance_locomotion(ANT_CFG, st, rng, False)


def make_humanoid(num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    st = _make_locomotion(HUMANOID_CFG, HUMANOID_DOF_LOWER, HUMANOID_DOF_UPPER, num_envs, seed,
                          reset_prob, True, fall_frac)
    st["motor_efforts"] = HUMANOID_MOTOR_EFFORTS.astype(F32)
    return st


def advance_humanoid(st, rng):
    _advance_locomotion(HUMANOID_CFG, st, rng, True)


# ---------------------------------------------------------------------------------------------
# Anymal (reference: tasks/anymal.py; cfg/task/Anymal.yaml) -- SURVEY.md 8f row 2
# ---------------------------------------------------------------------------------------------
ANYMAL_DT = 0.02
ANYMAL_CFG = dict(num_obs=48, num_actions=12, dt=ANYMAL_DT,
           max_episode_length=int(50 / ANYMAL_DT + 0.5), clip_obs=5.0, clip_actions=1.0,
           lin_vel_scale=2.0, ang_vel_scale=0.25, dof_pos_scale=1.0, dof_vel_scale=0.05, action_scale=0.5,
           rew_lin_vel_xy=1.0, rew_ang_vel_z=0.5, rew_torque=-0.000025,
           command_x_range=[-2.0, 2.0], command_y_range=[-1.0, 1.0], command_yaw_range=[-1.0, 1.0],
           num_bodies=17, base_index=0, knee_indices=[2, 6, 10, 14],
           # cfg/task/Anymal.yaml defaultJointAngles in DOF order (LF, LH, RF, RH: HAA, HFE, KFE)
           default_dof_pos=[0.03, 0.4, -0.8, 0.03, -0.4, 0.8, -0.03, 0.4, -0.8, -0.03, -0.4, 0.8])


def make_anymal(num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    n = num_envs
    g = np.random.Generator(np.random.PCG64(seed))
    f = lambda *s: g.standard_normal(s).astype(np.float32)
    root = np.zeros((n, 13), np.float32)
    root[:, :3] = f(n, 3)
    q = f(n, 4)
    half = n // 2
    q[:half] = np.array([0, 0, 0, 1], np.float32) + 0.15 * f(half, 4)       # mostly upright
    root[:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
    root[:, 7:13] = f(n, 6)
    contact = 0.4 * f(n, ANYMAL_CFG["num_bodies"], 3)                              # norms straddle the > 1 threshold
    progress = g.integers(0, ANYMAL_CFG["max_episode_length"] + 2, n).astype(np.int64)
    progress[: max(1, n // 16)] = ANYMAL_CFG["max_episode_length"] - 2             # time-outs this step
    return {
        "root_states": root, "dof_state": f(n, 12, 2) * np.array([0.5, 2.0], np.float32),
        "contact_forces": contact.astype(np.float32), "torques": 20.0 * f(n, 12),
        "actions": g.uniform(-1.5, 1.5, (n, 12)).astype(np.float32),
        "commands": g.uniform(-1, 1, (n, 3)).astype(np.float32),
        "reset_buf": (g.random(n) < reset_prob).astype(np.int64), "progress_buf": progress,
    }


def advance_anymal(st, g):
    for k in ("root_states", "dof_state", "contact_forces", "torques"):
        st[k] = (st[k] + 0.1 * g.standard_normal(st[k].shape)).astype(np.float32)
    q = st["root_states"][:, 3:7]
    st["root_states"][:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
    st["actions"] = g.uniform(-1.5, 1.5, st["actions"].shape).astype(np.float32)


# ---------------------------------------------------------------------------------------------
# BallBalance / Quadcopter / Ingenuity (reference: tasks/ball_balance.py, quadcopter.py, ingenuity.py and their
# cfg/task/*.yaml) -- SURVEY.md 8f row 2.  These reset in PRE-physics position: a state holds what
# pre_physics_step reads (flags, targets, thrusts) and the post-simulate tensors post_physics_step reads.
# ---------------------------------------------------------------------------------------------
BALL_BALANCE_CFG = dict(num_obs=24, num_actions=3, dt=0.01, action_speed_scale=20, max_episode_length=500,
                        clip_obs=5.0, clip_actions=1.0, ball_radius=0.1,
                        # generated MJCF (ball_balance.py:195-214): upper leg -45..45 deg, lower leg -70..90 deg
                        dof_lower=[-45 * _DEG, -70 * _DEG] * 3, dof_upper=[45 * _DEG, 90 * _DEG] * 3)
QUADCOPTER_CFG = dict(num_obs=21, num_actions=12, dt=0.01, max_episode_length=500, clip_obs=5.0, clip_actions=1.0,
                      num_bodies=9, rotor_bodies=[2, 4, 6, 8], max_thrust=2.0,
                      dof_lower=[-30 * _DEG] * 8, dof_upper=[30 * _DEG] * 8)      # quadcopter.py:180,199
INGENUITY_CFG = dict(num_obs=13, num_actions=6, dt=0.01, max_episode_length=2000, clip_obs=5.0, clip_actions=1.0,
                     num_bodies=6, thrust_upper_limit=2000.0, thrust_lateral_component=0.2)   # ingenuity.py:90-92


def _root_rows(g, n, pos_scale, z_lo, z_hi):
    root = np.zeros((n, 13), F32)
    root[:, 0:2] = pos_scale * g.standard_normal((n, 2))
    root[:, 2] = g.uniform(z_lo, z_hi, n)
    root[:, 3:7] = _unit_quats(g, n)
    root[:, 7:13] = g.standard_normal((n, 6))
    return root.astype(F32)


def make_ball_balance(num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    n, g, cfg = num_envs, _rng(seed), BALL_BALANCE_CFG
    root = np.zeros((n, 2, 13), F32)
    root[:, 0, 2], root[:, 0, 6] = 0.62, 1.0                                        # tray pose
    root[:, 1] = _root_rows(g, n, 0.3, 0.05 if fall_frac is None else 0.2 - 0.5 * fall_frac, 2.0)
    init_root = root.copy()
    init_root[:, 1] = _root_rows(g, n, 0.1, 0.9, 1.1)
    dof = (g.standard_normal((n, 6, 2)) * np.array([0.4, 2.0])).astype(F32)
    lo, hi = np.array(cfg["dof_lower"]), np.array(cfg["dof_upper"])
    return {
        "root_states": root, "dof_states": dof, "sensor_tensor": (30.0 * g.standard_normal((n, 3, 6))).astype(F32),
        "initial_root_states": init_root, "initial_dof_states": (0.05 * g.standard_normal((n, 6, 2))).astype(F32),
        "dof_position_targets": g.uniform(1.2 * lo, 1.2 * hi, (n, 6)).astype(F32),   # some beyond the limits
        "actions": g.uniform(-1.5, 1.5, (n, 3)).astype(F32),
        "reset_buf": _resets(g, n, reset_prob), "progress_buf": _progress(g, n, cfg["max_episode_length"]),
    }


def advance_ball_balance(st, g):
    n = st["reset_buf"].shape[0]
    st["root_states"][:, 1] = _root_rows(g, n, 0.3, 0.05, 2.0)
    st["dof_states"] = (st["dof_states"] + 0.1 * g.standard_normal(st["dof_states"].shape)).astype(F32)
    st["sensor_tensor"] = (30.0 * g.standard_normal(st["sensor_tensor"].shape)).astype(F32)
    st["actions"] = g.uniform(-1.5, 1.5, st["actions"].shape).astype(F32)


def make_quadcopter(num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    n, g, cfg = num_envs, _rng(seed), QUADCOPTER_CFG
    init_root = np.zeros((n, 13), F32)
    init_root[:, 2], init_root[:, 6] = 1.0, 1.0
    forces = np.zeros((n, cfg["num_bodies"], 3), F32)
    forces[:, cfg["rotor_bodies"], 2] = g.uniform(0, 2, (n, 4))
    return {
        "root_states": _root_rows(g, n, 1.2, 0.1, 2.5), "dof_states": (g.standard_normal((n, 8, 2)) * np.array([0.3, 2.0])).astype(F32),
        "initial_root_states": init_root, "initial_dof_states": np.zeros((n, 8, 2), F32),
        "dof_position_targets": g.uniform(-0.7, 0.7, (n, 8)).astype(F32), "thrusts": g.uniform(-0.2, 2.2, (n, 4)).astype(F32),
        "forces": forces, "actions": g.uniform(-1.5, 1.5, (n, 12)).astype(F32),
        "reset_buf": _resets(g, n, reset_prob), "progress_buf": _progress(g, n, cfg["max_episode_length"]),
    }


def advance_quadcopter(st, g):
    n = st["reset_buf"].shape[0]
    st["root_states"] = _root_rows(g, n, 1.2, 0.1, 2.5)
    st["dof_states"] = (st["dof_states"] + 0.1 * g.standard_normal(st["dof_states"].shape)).astype(F32)
    st["actions"] = g.uniform(-1.5, 1.5, st["actions"].shape).astype(F32)


def make_ingenuity(num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    n, g, cfg = num_envs, _rng(seed), INGENUITY_CFG
    root = np.zeros((n, 2, 13), F32)
    root[:, 0] = _root_rows(g, n, 3.0, 0.2, 3.0)
    root[:, 1, 6] = 1.0
    tgt = np.concatenate([g.uniform(-5, 5, (n, 2)), g.uniform(1, 2, (n, 1))], 1).astype(F32)
    root[:, 1, 0:3] = tgt + np.array([0, 0, 0.4], F32)
    init_root = np.zeros((n, 13), F32)
    init_root[:, 2], init_root[:, 6] = 1.0, 1.0
    progress = _progress(g, n, cfg["max_episode_length"])
    retarget = g.random(n) < 0.05
    progress[retarget] = 500 * g.integers(0, 4, int(retarget.sum()))             # progress % 500 == 0 rows (ingenuity.py:325)
    return {
        "root_tensor": root, "dof_states": (g.standard_normal((n, 4, 2)) * np.array([0.3, 30.0])).astype(F32),
        "initial_root_states": init_root, "target_root_positions": tgt,
        "thrusts": g.standard_normal((n, 2, 3)).astype(F32), "forces": g.standard_normal((n, cfg["num_bodies"], 3)).astype(F32),
        "actions": g.uniform(-1.5, 1.5, (n, 6)).astype(F32),
        "reset_buf": _resets(g, n, reset_prob), "progress_buf": progress,
    }


def advance_ingenuity(st, g):
    n = st["reset_buf"].shape[0]
    st["root_tensor"][:, 0] = _root_rows(g, n, 3.0, 0.2, 3.0)
    st["dof_states"] = (st["dof_states"] + 0.1 * g.standard_normal(st["dof_states"].shape)).astype(F32)
    st["actions"] = g.uniform(-1.5, 1.5, st["actions"].shape).astype(F32)


MAKERS = {"cartpole": make_cartpole, "ant": make_ant, "humanoid": make_humanoid, "anymal": make_anymal,
          "ball_balance": make_ball_balance, "quadcopter": make_quadcopter, "ingenuity": make_ingenuity}
ADVANCERS = {"cartpole": advance_cartpole, "ant": advance_ant, "humanoid": advance_humanoid,
             "anymal": advance_anymal, "ball_balance": advance_ball_balance, "quadcopter": advance_quadcopter,
             "ingenuity": advance_ingenuity}


def make_state(task, num_envs, seed=0, reset_prob=0.02, **kw):
    return MAKERS[task](num_envs, seed=seed, reset_prob=reset_prob, **kw)


def advance_physics(task, st, rng):
    ADVANCERS[task](st, rng)


# ---------------------------------------------------------------------------------------------
# ShadowHand (reference: tasks/shadow_hand.py; cfg/task/ShadowHand.yaml)
# ---------------------------------------------------------------------------------------------
SHADOW_HAND_CFG = dict(
    max_episode_length=600, dist_reward_scale=-10.0, rot_reward_scale=1.0, rot_eps=0.1,
    action_penalty_scale=-0.0002, reach_goal_bonus=250.0, fall_dist=0.24, fall_penalty=0.0,
    success_tolerance=0.1, max_consecutive_successes=0, av_factor=0.1, vel_obs_scale=0.2,
    force_torque_obs_scale=10.0, reset_position_noise=0.01, reset_dof_pos_noise=0.2,
    reset_dof_vel_noise=0.0, clip_obs=5.0, clip_actions=1.0, dt=0.0166, ignore_z_rot=False,
    use_relative_control=False, dof_speed_scale=20.0, act_moving_average=1.0,
    force_prob_range=(0.001, 0.1), goal_displacement=(-0.2, -0.06, 0.12),
    num_hand_dofs=24, num_actions=20, num_fingertips=5, num_bodies=28,
    fingertip_handles=(7, 11, 15, 20, 25),
    obs_type="full_state", asymmetric_obs=False)
SHADOW_HAND_NUM_OBS = {"openai": 42, "full_no_vel": 77, "full": 157, "full_state": 211}
# AllegroHand (tasks/allegro_hand.py; cfg/task/AllegroHand.yaml): ShadowHand's reset / reward code on a 16-DOF hand with all
# joints actuated, no fingertip or force-sensor observations (allegro_hand.py:108-126, :439-489)
ALLEGRO_HAND_CFG = dict(SHADOW_HAND_CFG, hand="allegro", num_hand_dofs=16, num_actions=16, num_fingertips=0, num_bodies=8,
                        fingertip_handles=(0, 0, 0, 0, 0), num_states_full=88)
ALLEGRO_HAND_NUM_OBS = {"full_no_vel": 50, "full": 72, "full_state": 88}
# the 20 actuated of the 24 hand DOFs (coupled distal joints are passive in the MJCF)
SHADOW_HAND_ACTUATED = (0, 1, 2, 3, 4, 6, 7, 8, 10, 11, 12, 14, 15, 16, 17, 19, 20, 21, 22, 23)


def _shadow_limits(d=24):
    r = _rng(424242)
    lo = r.uniform(-0.8, 0.0, 24).astype(F32)
    hi = r.uniform(0.5, 1.6, 24).astype(F32)
    return lo[:d].copy(), hi[:d].copy()


def _quat_mul_np(a, b):
    x1, y1, z1, w1 = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    x2, y2, z2, w2 = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                     w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2], -1)


def _rand_unit_quats(rng, n):
    q = rng.standard_normal((n, 4))
    return q / np.linalg.norm(q, axis=-1, keepdims=True)


def _shadow_sim_draw(st, rng, cfg):
    """(Re)draw everything the simulator owns for one step."""
    n, d, b = st["progress_buf"].shape[0], cfg["num_hand_dofs"], cfg["num_bodies"]
    lo, hi = st["shadow_hand_dof_lower_limits"], st["shadow_hand_dof_upper_limits"]
    span = (hi - lo).astype(np.float64)
    dof = np.zeros((n, d, 2), F32)
    dof[:, :, 0] = rng.uniform(lo - 0.05 * span, hi + 0.05 * span, (n, d))
    dof[:, :, 1] = rng.normal(0.0, 2.0, (n, d))
    st["dof_state"] = dof.reshape(n * d, 2)
    root = st["root_state_tensor"].reshape(n, 3, 13)
    goal = st["goal_states"]
    # object near the goal: straddles fall_dist = 0.24 and, for 5 % of envs, the success tolerance
    root[:, 1, 0:3] = goal[:, 0:3] + rng.normal(0.0, 0.1, (n, 3))
    q = _rand_unit_quats(rng, n)
    close = rng.random(n) < 0.05
    ang = rng.uniform(-0.04, 0.04, n)
    small = np.stack([np.sin(ang / 2), np.zeros(n), np.zeros(n), np.cos(ang / 2)], -1)
    q[close] = _quat_mul_np(goal[close, 3:7].astype(np.float64), small[close])
    root[:, 1, 3:7] = q
    root[:, 1, 7:13] = rng.standard_normal((n, 6))
    root[:, 0, 7:13] = 0.0
    rb = np.zeros((n, b, 13), F32)
    rb[:, :, 0:3] = rng.normal(0.0, 0.3, (n, b, 3))
    rb[:, :, 3:7] = _rand_unit_quats(rng, n * b).reshape(n, b, 4)
    rb[:, :, 7:13] = rng.standard_normal((n, b, 6))
    st["rigid_body_states"] = rb
    st["vec_sensor_tensor"] = rng.normal(0.0, 0.3, (n, max(cfg["num_fingertips"], 1) * 6)).astype(F32)
    st["dof_force_tensor"] = rng.normal(0.0, 0.3, (n, d)).astype(F32)
    st["actions"] = rng.uniform(-1.0, 1.0, (n, cfg["num_actions"])).astype(F32)


def make_shadow_hand(num_envs, seed=0, reset_prob=0.02, fall_frac=None, obs_type=None, asymmetric_obs=None,
                     max_consecutive_successes=None, goal_reset_prob=0.05, base_cfg=None):
    cfg = dict(base_cfg or SHADOW_HAND_CFG)
    allegro = cfg.get("hand") == "allegro"
    if obs_type is not None:
        cfg["obs_type"] = obs_type
    if asymmetric_obs is not None:
        cfg["asymmetric_obs"] = asymmetric_obs
    if max_consecutive_successes is not None:
        cfg["max_consecutive_successes"] = max_consecutive_successes
    rng = _rng(seed)
    n, d = num_envs, cfg["num_hand_dofs"]
    st = {"cfg_obs_type": np.array(cfg["obs_type"]), "cfg_asymmetric_obs": np.array(int(cfg["asymmetric_obs"])),
          "cfg_max_consecutive_successes": np.array(int(cfg["max_consecutive_successes"]))}
    lo, hi = _shadow_limits(d)
    st["shadow_hand_dof_lower_limits"], st["shadow_hand_dof_upper_limits"] = lo, hi
    st["shadow_hand_dof_default_pos"] = np.zeros(d, F32)
    st["shadow_hand_dof_default_vel"] = np.zeros(d, F32)
    obj_init = np.zeros((n, 13), F32)
    obj_init[:, 0:3] = np.array([0.0, -0.39, 0.54], F32) + rng.normal(0, 0.5, (n, 3)).astype(F32)
    obj_init[:, 6] = 1.0
    st["object_init_state"] = obj_init
    goal = obj_init.copy()
    goal[:, 2] -= F32(0.04)                       # shadow_hand.py:402-404
    st["goal_init_state"] = goal.copy()
    goal[:, 3:7] = _rand_unit_quats(rng, n)
    st["goal_states"] = goal
    root = np.zeros((n, 3, 13), F32)
    root[:, 0, 0:3] = obj_init[:, 0:3] + np.array([0.0, 0.39, -0.04], F32)
    root[:, 0, 6] = 1.0
    root[:, 2, 0:3] = goal[:, 0:3] + np.array(cfg["goal_displacement"], F32)
    root[:, 2, 3:7] = goal[:, 3:7]
    st["root_state_tensor"] = root.reshape(n * 3, 13)
    st["hand_indices"] = (3 * np.arange(n)).astype(I64)
    st["object_indices"] = (3 * np.arange(n) + 1).astype(I64)
    st["goal_object_indices"] = (3 * np.arange(n) + 2).astype(I64)
    st["fingertip_handles"] = np.array(cfg["fingertip_handles"], I64)
    st["actuated_dof_indices"] = np.arange(d, dtype=I64) if allegro else np.array(SHADOW_HAND_ACTUATED, I64)
    st["progress_buf"] = _progress(rng, n, cfg["max_episode_length"])
    st["reset_buf"] = _resets(rng, n, reset_prob)
    st["reset_goal_buf"] = _resets(rng, n, goal_reset_prob if reset_prob < 1.0 else 1.0)
    st["randomize_buf"] = np.zeros(n, I64)
    st["successes"] = rng.integers(0, 6, n).astype(F32)
    st["consecutive_successes"] = np.array([1.5], F32)
    st["prev_targets"] = rng.uniform(lo, hi, (n, d)).astype(F32)
    st["cur_targets"] = st["prev_targets"].copy()
    st["rb_forces"] = rng.normal(0, 1, (n, cfg["num_bodies"], 3)).astype(F32)
    st["random_force_prob"] = rng.uniform(0.001, 0.1, n).astype(F32)
    _shadow_sim_draw(st, rng, cfg)
    nobs = (ALLEGRO_HAND_NUM_OBS if allegro else SHADOW_HAND_NUM_OBS)[cfg["obs_type"]]
    st["obs_buf"] = np.zeros((n, nobs), F32)
    st["states_buf"] = np.zeros((n, (88 if allegro else 211) if cfg["asymmetric_obs"] else 0), F32)
    st["rew_buf"] = np.zeros(n, F32)
    return st


def advance_shadow_hand(st, rng):
    _shadow_sim_draw(st, rng, SHADOW_HAND_CFG)


def make_allegro_hand(num_envs, seed=0, reset_prob=0.02, **kw):
    return make_shadow_hand(num_envs, seed=seed, reset_prob=reset_prob, base_cfg=ALLEGRO_HAND_CFG, **kw)


def advance_allegro_hand(st, rng):
    _shadow_sim_draw(st, rng, ALLEGRO_HAND_CFG)


MAKERS["shadow_hand"] = make_shadow_hand
ADVANCERS["shadow_hand"] = advance_shadow_hand
MAKERS["allegro_hand"] = make_allegro_hand
ADVANCERS["allegro_hand"] = advance_allegro_hand


# ---------------------------------------------------------------------------------------------
# Franka multi-agent tasks (reference: tasks/franka_{reach,collect,combine}_MA.py, franka_PPMA.py)
# ---------------------------------------------------------------------------------------------
FRANKA_VARIANTS = ("reach", "collect", "combine", "pp")
FRANKA_DEFAULT_DOF_POS = np.array([0, 0.1963, 0, -2.6180, 0, 2.9416, 0.7854, 0.035, 0.035], F32)
# franka_panda_gripper.urdf joint limits (7 arm joints + 2 fingers)
FRANKA_DOF_LOWER = np.array([-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973, 0.0, 0.0], F32)
FRANKA_DOF_UPPER = np.array([2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973, 0.04, 0.04], F32)
FRANKA_LINKS = 17          # rigid bodies per arm (franka_panda_gripper.urdf)
FRANKA_LINK_BASE, FRANKA_LINK_HAND, FRANKA_LINK_GRIP_SITE = 0, 9, 14


def franka_ma_cfg(variant, num_agents=2, num_targets=None):
    """Layout + parameters of one Franka MA variant.  Actor / body order (one env):
    table, stand, [wall | stack_base], cubes x T, frankas x A (17 links each), [dests x T]."""
    a = num_agents
    t = a if num_targets is None else num_targets
    extra_front = 1 if variant in ("collect", "combine") else 0
    extra_back = t if variant == "pp" else 0
    cube0 = 2 + extra_front
    franka0 = cube0 + t
    body_franka0 = cube0 + t
    num_obs = {"reach": 10 + 3 * t + 3 * (a - 1), "collect": 3 * t + 17 + 3 * (a - 1) + 2,
               "combine": 7 * t + 3 + 7 * a + 13 + 2 + 1 + 1, "pp": 7 * t + 3 * t + 7 * a + 13 + 2 + 1}[variant]
    return dict(
        variant=variant, num_agents=a, num_targets=t,
        num_actions=6 if variant == "reach" else 7, num_obs=num_obs,
        max_episode_length=150 if variant == "reach" else 300, clip_obs=5.0, clip_actions=1.0,
        start_position_noise=0.25, franka_dof_noise=0.25, cubeA_size=0.05, dest_size=0.05,
        table_surface_pos=(0.0, 0.0, 1.025), wall_pos_y=0.3, wall_height=0.3,
        rand_z_range={"reach": 0.5, "collect": 0.2, "combine": 0.2, "pp": 0.2}[variant],
        num_actors=2 + extra_front + t + a + extra_back, num_bodies=2 + extra_front + t + FRANKA_LINKS * a + extra_back,
        num_dofs=9 * a,
        cube_actor_start=cube0, cube_actor_step=1,
        franka_actor_start=franka0,
        dest_actor_start=franka0 + a, dest_actor_step=1,
        stack_base_actor=2,
        eef_body_start=body_franka0 + FRANKA_LINK_GRIP_SITE, eef_body_step=FRANKA_LINKS,
        base_body_start=body_franka0 + FRANKA_LINK_BASE, base_body_step=FRANKA_LINKS,
        hand_body_start=body_franka0 + FRANKA_LINK_HAND, hand_body_step=FRANKA_LINKS)


def _franka_sim_draw(st, rng, cfg):
    n, a, t = st["progress_buf"].shape[0] // cfg["num_agents"], cfg["num_agents"], cfg["num_targets"]
    nb, na = cfg["num_bodies"], cfg["num_actors"]
    v = cfg["variant"]
    rb = np.zeros((n, nb, 13), F32)
    rb[:, :, 0] = rng.uniform(-0.7, 0.7, (n, nb))
    rb[:, :, 1] = rng.uniform(-0.5, 0.8, (n, nb))
    rb[:, :, 2] = rng.uniform(1.0, 1.6, (n, nb))
    rb[:, :, 3:7] = _rand_unit_quats(rng, n * nb).reshape(n, nb, 4)
    rb[:, :, 7:13] = rng.standard_normal((n, nb, 6))
    root = np.zeros((n, na, 13), F32)
    root[:, :, 6] = 1.0
    cs, fs = cfg["cube_actor_start"], cfg["franka_actor_start"]
    cubes = root[:, cs:cs + t]
    if v == "collect":     # straddle every FSM threshold of franka_collect_MA.py:549-606
        cubes[:, :, 0] = rng.uniform(-0.8, 0.8, (n, t))
        cubes[:, :, 1] = rng.uniform(-0.3, 0.7, (n, t))
        cubes[:, :, 2] = rng.uniform(1.0, 1.6, (n, t))
    else:
        cubes[:, :, 0:2] = rng.uniform(-0.25, 0.25, (n, t, 2))
        cubes[:, :, 2] = 1.05 + rng.uniform(0.0, cfg["rand_z_range"], (n, t))
    cubes[:, :, 3:7] = _rand_unit_quats(rng, n * t).reshape(n, t, 4)
    cubes[:, :, 7:13] = rng.normal(0, 0.2, (n, t, 6))
    if v == "pp":
        ds = cfg["dest_actor_start"]
        dests = root[:, ds:ds + t]
        dests[:, :, 0] = np.array([-0.25, 0.25] * (t // 2 + 1), F32)[:t] + rng.uniform(-0.25, 0.25, (n, t))
        dests[:, :, 1] = 0.3 + rng.uniform(-0.25, 0.25, (n, t))
        dests[:, :, 2] = 1.01875
        near = rng.random((n, t)) < 0.4     # cube hovering over / next to its destination (states 3..6)
        off = np.zeros((n, t, 3), F32)
        off[:, :, 0:2] = rng.uniform(-0.04, 0.04, (n, t, 2))
        off[:, :, 2] = rng.uniform(-0.12, 0.12, (n, t))
        dests[near, 0:3] = (cubes[:, :, 0:3] + off)[near]
    if v == "combine":
        sb = root[:, cfg["stack_base_actor"]]
        sb[:, 0:2] = rng.uniform(-0.15, 0.15, (n, 2))
        sb[:, 2] = 1.01875 + rng.uniform(-0.05, 0.08, n)
    # a fraction of the grippers sits on a cube (own cube for combine/pp, any cube otherwise)
    es, est = cfg["eef_body_start"], cfg["eef_body_step"]
    for ag in range(a):
        on = rng.random(n) < 0.3
        tgt = cubes[:, ag % t, 0:3] + rng.uniform(-0.02, 0.02, (n, 3)).astype(F32)
        rb[on, es + ag * est, 0:3] = tgt[on]
    # two grippers of an env sometimes too close (gFSM of combine / pp, franka_PPMA.py:597-611)
    if a >= 2:
        close = rng.random(n) < 0.15
        rb[close, es + est, 0:3] = rb[close, es, 0:3] + rng.uniform(-0.1, 0.1, (int(close.sum()), 3)).astype(F32)
    st["_rigid_body_state"] = rb
    st["_root_state"] = root
    st["_contact_forces"] = rng.normal(0.0, 0.1, (n, nb, 3)).astype(F32)
    dof = np.zeros((n, cfg["num_dofs"], 2), F32)
    dof[:, :, 0] = rng.uniform(np.tile(FRANKA_DOF_LOWER, a), np.tile(FRANKA_DOF_UPPER, a), (n, cfg["num_dofs"]))
    dof[:, :, 1] = rng.normal(0, 1, (n, cfg["num_dofs"]))
    st["_dof_state"] = dof
    st["actions"] = rng.uniform(-1.0, 1.0, (n * a, cfg["num_actions"])).astype(F32)


def make_franka_ma(variant, num_envs, seed=0, reset_prob=0.02, fall_frac=None, num_agents=2, num_targets=None,
                   ragged_flags=False):
    cfg = franka_ma_cfg(variant, num_agents, num_targets)
    rng = _rng(seed)
    n, a, t = num_envs, cfg["num_agents"], cfg["num_targets"]
    st = {"cfg_variant": np.array(variant), "cfg_num_agents": np.array(a), "cfg_num_targets": np.array(t)}
    env_prog = _progress(rng, n, cfg["max_episode_length"])
    st["progress_buf"] = np.repeat(env_prog, a)           # lock-step across an env's agents
    env_reset = _resets(rng, n, reset_prob)
    st["reset_buf"] = np.repeat(env_reset, a)
    if ragged_flags:                                      # exercise the AND filter (franka_reach_MA.py:879-882)
        flip = rng.random(n * a) < 0.2
        st["reset_buf"] = np.where(flip, 1 - st["reset_buf"], st["reset_buf"]).astype(I64)
    _franka_sim_draw(st, rng, cfg)
    st["_pos_control"] = rng.normal(0, 1, (n, cfg["num_dofs"])).astype(F32)
    st["_effort_control"] = rng.normal(0, 1, (n, cfg["num_dofs"])).astype(F32)
    st["_init_cubeA_state"] = np.zeros((n, t, 13), F32)
    st["obs_buf"] = np.zeros((n * a, cfg["num_obs"]), F32)
    st["rew_buf"] = np.zeros(n * a, F32)
    return st


def _advance_franka(st, rng):
    cfg = franka_ma_cfg(str(st["cfg_variant"]), int(st["cfg_num_agents"]), int(st["cfg_num_targets"]))
    _franka_sim_draw(st, rng, cfg)


for _v in FRANKA_VARIANTS:
    MAKERS["franka_" + _v] = (lambda v: (lambda num_envs, seed=0, reset_prob=0.02, **kw: make_franka_ma(
        v, num_envs, seed=seed, reset_prob=reset_prob, **kw)))(_v)
    ADVANCERS["franka_" + _v] = _advance_franka


# ---------------------------------------------------------------------------------------------
# Single-agent Franka tasks (reference: tasks/franka_reach.py, franka_cube_stack2.py,
# franka_cube_stack.py, franka_cabinet.py and their cfg/task/*.yaml)
# ---------------------------------------------------------------------------------------------
FRANKA_SA_VARIANTS = ("reach", "stack2", "stack")
FRANKA_SA_NUM_OBS = {"reach": 14, "stack2": 21, "stack": 19}


def franka_sa_cfg(variant):
    """Layout + parameters of FrankaReach / FrankaCubeStack2 / FrankaCubeStack.  Actor order of one env: franka
    (17 links), table, table stand, cubeA[, cubeB] (franka_reach.py:300-330, franka_cube_stack2.py:303-330)."""
    two = variant != "reach"
    return dict(
        variant=variant, num_obs=FRANKA_SA_NUM_OBS[variant], num_actions=7,
        max_episode_length=300 if variant == "stack" else 150, clip_obs=5.0, clip_actions=1.0,
        start_position_noise=0.25, start_rotation_noise=0.785, franka_dof_noise=0.25,
        cubeA_size=0.05, cubeB_size=0.07, table_surface_pos=(0.0, 0.0, 1.025),
        rand_z_range=0.0 if variant == "stack" else 0.5,
        r_dist_scale=0.1, r_lift_scale=1.5, r_align_scale=2.0, r_stack_scale=16.0,
        num_actors=5 if two else 4, num_bodies=FRANKA_LINKS + (4 if two else 3), num_dofs=9,
        eef_body=FRANKA_LINK_GRIP_SITE, lf_body=15, rf_body=16, cubeA_actor=3, cubeB_actor=4 if two else 3)


def _franka_sa_sim_draw(st, g, cfg):
    n, v = st["progress_buf"].shape[0], cfg["variant"]
    nb, na = cfg["num_bodies"], cfg["num_actors"]
    rb = np.zeros((n, nb, 13), F32)
    rb[:, :, 0:2] = g.uniform(-0.5, 0.5, (n, nb, 2))
    rb[:, :, 2] = g.uniform(1.0, 1.6, (n, nb))
    rb[:, :, 3:7] = _rand_unit_quats(g, n * nb).reshape(n, nb, 4)
    rb[:, :, 7:13] = g.standard_normal((n, nb, 6))
    root = np.zeros((n, na, 13), F32)
    root[:, :, 6] = 1.0
    A = root[:, cfg["cubeA_actor"]]
    A[:, 0:2] = g.uniform(-0.25, 0.25, (n, 2))
    A[:, 2] = 1.05 + g.uniform(0.0, 0.5, n)
    A[:, 3:7] = _rand_unit_quats(g, n)
    A[:, 7:13] = g.normal(0, 0.2, (n, 6))
    if v != "reach":
        B = root[:, cfg["cubeB_actor"]]
        B[:, 0:2] = g.uniform(-0.25, 0.25, (n, 2))
        B[:, 2] = 1.06 + g.uniform(0.0, 0.05, n)
        B[:, 3:7] = _rand_unit_quats(g, n)
        # cube A hovering over / resting on cube B: FSM states 4-6 of Stack2, the stack success of Stack
        near = g.random(n) < 0.45
        off = np.zeros((n, 3), F32)
        off[:, 0:2] = g.uniform(-0.03, 0.03, (n, 2))
        off[:, 2] = g.uniform(0.03, 0.12, n)
        exact = g.random(n) < 0.4
        off[exact, 0:2] = g.uniform(-0.012, 0.012, (int(exact.sum()), 2))
        off[exact, 2] = 0.06 + g.uniform(-0.004, 0.004, int(exact.sum()))
        A[near, 0:3] = (B[:, 0:3] + off)[near]
        rest = near & exact & (g.random(n) < 0.5)          # cube B on the table, A on top: |height - target| < 0.02
        B[rest, 2] = 1.025 + 0.035
        A[rest, 2] = 1.025 + 0.07 + 0.025 + g.uniform(-0.01, 0.01, int(rest.sum()))
    on = g.random(n) < 0.4                                 # gripper on the cube: FSM >= 1
    rb[on, cfg["eef_body"], 0:3] = (A[:, 0:3] + g.uniform(-0.018, 0.018, (n, 3)).astype(F32))[on]
    for b in (cfg["lf_body"], cfg["rf_body"]):
        rb[:, b, 0:3] = rb[:, cfg["eef_body"], 0:3] + g.uniform(-0.05, 0.05, (n, 3)).astype(F32)
    st["rigid_body_state"], st["root_state"] = rb, root
    dof = np.zeros((n, 9, 2), F32)
    dof[:, :, 0] = g.uniform(FRANKA_DOF_LOWER, FRANKA_DOF_UPPER, (n, 9))
    dof[:, :, 1] = g.normal(0, 1, (n, 9))
    st["dof_state"] = dof
    st["actions"] = g.uniform(-1.3, 1.3, (n, 7)).astype(F32)


def make_franka_sa(variant, num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    cfg, g, n = franka_sa_cfg(variant), _rng(seed), num_envs
    st = {"cfg_variant": np.array(variant), "progress_buf": _progress(g, n, cfg["max_episode_length"]),
          "reset_buf": _resets(g, n, reset_prob)}
    _franka_sa_sim_draw(st, g, cfg)
    st["pos_control"] = g.normal(0, 1, (n, 9)).astype(F32)
    st["effort_control"] = g.normal(0, 1, (n, 9)).astype(F32)
    st["init_cubeA_state"] = np.zeros((n, 13), F32)
    st["init_cubeB_state"] = np.zeros((n, 13), F32)
    return st


def _advance_franka_sa(st, g):
    _franka_sa_sim_draw(st, g, franka_sa_cfg(str(st["cfg_variant"])))


for _v in FRANKA_SA_VARIANTS:
    MAKERS["franka_sa_" + _v] = (lambda v: (lambda num_envs, seed=0, reset_prob=0.02, **kw: make_franka_sa(
        v, num_envs, seed=seed, reset_prob=reset_prob, **kw)))(_v)
    ADVANCERS["franka_sa_" + _v] = _advance_franka_sa

# franka_panda.urdf (11 bodies, 9 DOFs) + sektion_cabinet_2.urdf (5 bodies, 4 DOFs) + props; franka_cabinet.py:139-330
FRANKA_CABINET_CFG = dict(
    num_obs=23, num_actions=9, max_episode_length=500, clip_obs=5.0, clip_actions=1.0, dt=0.0166,
    action_scale=7.5, dof_vel_scale=0.1, dist_reward_scale=2.0, rot_reward_scale=0.5, around_handle_reward_scale=0.25,
    open_reward_scale=7.5, finger_dist_reward_scale=5.0, action_penalty_scale=0.01, distX_offset=0.04,
    num_props=16, num_dofs=13, num_bodies=11 + 5 + 16, num_actors=2 + 16,
    hand_body=8, lfinger_body=9, rfinger_body=10, drawer_body=11 + 3,
    default_dof_pos=[1.157, -1.066, -0.155, -2.239, -1.841, 1.003, 0.469, 0.035, 0.035],   # franka_cabinet.py:110
    dof_speed_scales=[1.0] * 7 + [0.1, 0.1],                                                # :253-254
    # compute_grasp_transforms inputs (:300-326): grasp pose between the finger pads, drawer handle offset
    franka_local_grasp_pos=[0.0, 0.04, 0.0584], franka_local_grasp_rot=[0.0, 0.0, 0.0, 1.0],
    drawer_local_grasp_pos=[0.3, 0.01, 0.0], drawer_local_grasp_rot=[0.0, 0.0, 0.0, 1.0],
    gripper_forward_axis=[0, 0, 1], drawer_inward_axis=[-1, 0, 0], gripper_up_axis=[0, 1, 0], drawer_up_axis=[0, 0, 1])


def _franka_cabinet_sim_draw(st, g):
    cfg, n = FRANKA_CABINET_CFG, st["progress_buf"].shape[0]
    nb = cfg["num_bodies"]
    rb = np.zeros((n, nb, 13), F32)
    rb[:, :, 0:3] = g.uniform(-0.6, 0.6, (n, nb, 3)) + np.array([0, 0, 0.6], F32)
    rb[:, :, 3:7] = _rand_unit_quats(g, n * nb).reshape(n, nb, 4)
    rb[:, :, 7:13] = g.standard_normal((n, nb, 6))
    # half of the hands sit at the drawer handle with a nearly aligned gripper (around-handle / finger terms)
    near = g.random(n) < 0.5
    k = int(near.sum())
    small = np.zeros((k, 4), F32)
    small[:, 0:3] = g.normal(0, 0.05, (k, 3))
    small[:, 3] = 1.0
    small /= np.linalg.norm(small, axis=1, keepdims=True)
    rb[near, cfg["drawer_body"], 3:7] = small
    rb[near, cfg["hand_body"], 0:3] = rb[near, cfg["drawer_body"], 0:3] + np.array([0.3, 0.0, -0.05], F32) + \
        g.normal(0, 0.03, (k, 3)).astype(F32)
    for b, dz in ((cfg["lfinger_body"], 0.03), (cfg["rfinger_body"], -0.03)):
        rb[near, b, 0:3] = rb[near, cfg["drawer_body"], 0:3] + np.array([0.3, 0.01, dz], F32) + \
            g.normal(0, 0.02, (k, 3)).astype(F32)
    st["rigid_body_states"] = rb
    root = np.zeros((n, cfg["num_actors"], 13), F32)
    root[:, :, 0:3] = g.uniform(-1, 1, (n, cfg["num_actors"], 3))
    root[:, :, 6] = 1.0
    st["root_state_tensor"] = root
    dof = np.zeros((n, cfg["num_dofs"], 2), F32)
    dof[:, :9, 0] = g.uniform(FRANKA_DOF_LOWER, FRANKA_DOF_UPPER, (n, 9))
    dof[:, 9:, 0] = g.uniform(0.0, 0.45, (n, 4))                  # drawer opening straddles 0.01 / 0.2 / 0.39
    dof[:, :, 1] = g.normal(0, 1, (n, cfg["num_dofs"]))
    st["dof_state"] = dof
    st["actions"] = g.uniform(-1.3, 1.3, (n, 9)).astype(F32)


def make_franka_cabinet(num_envs, seed=0, reset_prob=0.02, fall_frac=None):
    cfg, g, n = FRANKA_CABINET_CFG, _rng(seed), num_envs
    st = {"progress_buf": _progress(g, n, cfg["max_episode_length"]), "reset_buf": _resets(g, n, reset_prob)}
    _franka_cabinet_sim_draw(st, g)
    st["default_prop_states"] = g.uniform(-1, 1, (n, cfg["num_props"], 13)).astype(F32)
    lo, hi = FRANKA_DOF_LOWER, FRANKA_DOF_UPPER
    tg = np.zeros((n, cfg["num_dofs"]), F32)
    tg[:, :9] = g.uniform(lo - 0.05, hi + 0.05, (n, 9))
    st["franka_dof_targets"] = tg
    return st


def _advance_franka_cabinet(st, g):
    _franka_cabinet_sim_draw(st, g)


MAKERS["franka_cabinet"] = make_franka_cabinet
ADVANCERS["franka_cabinet"] = _advance_franka_cabinet
