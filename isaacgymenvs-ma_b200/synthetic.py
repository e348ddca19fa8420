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


MAKERS = {"cartpole": make_cartpole, "ant": make_ant, "humanoid": make_humanoid}
ADVANCERS = {"cartpole": advance_cartpole, "ant": advance_ant, "humanoid": advance_humanoid}


def make_state(task, num_envs, seed=0, reset_prob=0.02, **kw):
    return MAKERS[task](num_envs, seed=seed, reset_prob=reset_prob, **kw)


def advance_physics(task, st, rng):
    ADVANCERS[task](st, rng)
