"""Golden fixtures for FrankaReach / FrankaCubeStack2 / FrankaCubeStack / FrankaCabinet: the
UNMODIFIED reference classes (/root/reference/isaacgymenvs/tasks/{franka_reach,franka_cube_stack2,franka_cube_stack,
franka_cabinet}.py) stepped through ``VecTask.step`` on synthetic state, built with ``Task.__new__`` + ``FakeGym``.

``gym.simulate`` is hooked: the hook snapshots what pre_physics_step wrote ("mid/", FrankaCabinet's DOF targets) and
loads the step's synthetic simulator state ("phys/") before ``post_physics_step`` runs.  The OSC pre_physics_step of
the three cube tasks needs PhysX Jacobians / mass matrices and is replaced by its first line (``self.actions =
actions.clone()``), as for the MA tasks.

Container-only (needs /root/reference):  python tests/golden/gen_golden_franka_sa.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import harness as H  # noqa: E402
import ref_harness  # noqa: E402
import ref_shim  # noqa: E402
import franka_sa_harness as FH  # noqa: E402
from gen_golden_small import HookGym  # noqa: E402

syn = H.pkg.synthetic
t = ref_harness._t
_CLS = {"franka_sa_reach": ("tasks.franka_reach", "FrankaReach"), "franka_sa_stack2": ("tasks.franka_cube_stack2", "FrankaCubeStack2"),
        "franka_sa_stack": ("tasks.franka_cube_stack", "FrankaCubeStack"), "franka_cabinet": ("tasks.franka_cabinet", "FrankaCabinet")}


def _new(task, cfg, n):
    mod, name = _CLS[task]
    cls = getattr(ref_shim.ref_module(mod), name)
    obj = cls.__new__(cls)
    ref_harness._common(obj, cfg, n, cfg["num_obs"], cfg["num_actions"])
    obj.gym = HookGym()
    obj.obs_buf, obj.rew_buf = torch.zeros(n, cfg["num_obs"]), torch.zeros(n)
    obj.randomize_buf = torch.zeros(n, dtype=torch.long)
    return obj


def build_sa(task, st):
    cfg, n = FH.cfg_of(task), st["reset_buf"].shape[0]
    obj = _new(task, cfg, n)
    obj.control_type = "osc"
    obj.start_position_noise, obj.start_rotation_noise = cfg["start_position_noise"], cfg["start_rotation_noise"]
    obj.franka_dof_noise = cfg["franka_dof_noise"]
    obj.cubeA_size, obj.cubeB_size = cfg["cubeA_size"], cfg["cubeB_size"]
    obj._table_surface_pos = np.array(cfg["table_surface_pos"])
    obj.reward_settings = {"r_dist_scale": cfg["r_dist_scale"], "r_lift_scale": cfg["r_lift_scale"],
                           "r_align_scale": cfg["r_align_scale"], "r_stack_scale": cfg["r_stack_scale"],
                           "table_height": obj._table_surface_pos[2]}
    obj.franka_default_dof_pos = t(syn.FRANKA_DEFAULT_DOF_POS)
    obj.franka_dof_lower_limits, obj.franka_dof_upper_limits = t(syn.FRANKA_DOF_LOWER), t(syn.FRANKA_DOF_UPPER)
    obj._rigid_body_state, obj._root_state, obj._dof_state = t(st["rigid_body_state"]), t(st["root_state"]), t(st["dof_state"])
    obj._q, obj._qd = obj._dof_state[..., 0], obj._dof_state[..., 1]
    obj._eef_state = obj._rigid_body_state[:, cfg["eef_body"], :]
    obj._eef_lf_state = obj._rigid_body_state[:, cfg["lf_body"], :]
    obj._eef_rf_state = obj._rigid_body_state[:, cfg["rf_body"], :]
    obj._cubeA_state = obj._root_state[:, cfg["cubeA_actor"], :]
    obj._cubeB_state = obj._root_state[:, cfg["cubeB_actor"], :]
    obj._init_cubeA_state, obj._init_cubeB_state = t(st["init_cubeA_state"]), t(st["init_cubeB_state"])
    obj._pos_control, obj._effort_control = t(st["pos_control"]), t(st["effort_control"])
    obj._global_indices = torch.arange(n * cfg["num_actors"], dtype=torch.int32).view(n, -1)
    obj.states = {"cubeA_size": torch.ones_like(obj._eef_state[:, 0]) * obj.cubeA_size,     # franka_reach.py:381
                  "cubeB_size": torch.ones_like(obj._eef_state[:, 0]) * obj.cubeB_size}
    obj.actions = None
    obj.pre_physics_step = lambda act: setattr(obj, "actions", act.clone())
    return obj


def build_cabinet(st):
    cfg, n = FH.cfg_of("franka_cabinet"), st["reset_buf"].shape[0]
    obj = _new("franka_cabinet", cfg, n)
    for k in ("dt", "action_scale", "dof_vel_scale", "dist_reward_scale", "rot_reward_scale", "around_handle_reward_scale",
              "open_reward_scale", "finger_dist_reward_scale", "action_penalty_scale", "distX_offset", "num_props"):
        setattr(obj, k, cfg[k])
    obj.num_franka_dofs = 9
    obj.hand_handle, obj.drawer_handle = cfg["hand_body"], cfg["drawer_body"]
    obj.lfinger_handle, obj.rfinger_handle = cfg["lfinger_body"], cfg["rfinger_body"]
    obj.rigid_body_states, obj.root_state_tensor, obj.dof_state = (t(st["rigid_body_states"]), t(st["root_state_tensor"]),
                                                                   t(st["dof_state"]))
    obj.franka_dof_state = obj.dof_state[:, :9]                                                  # franka_cabinet.py:99-107
    obj.franka_dof_pos, obj.franka_dof_vel = obj.franka_dof_state[..., 0], obj.franka_dof_state[..., 1]
    obj.cabinet_dof_state = obj.dof_state[:, 9:]
    obj.cabinet_dof_pos, obj.cabinet_dof_vel = obj.cabinet_dof_state[..., 0], obj.cabinet_dof_state[..., 1]
    obj.prop_states = obj.root_state_tensor[:, 2:]
    obj.default_prop_states = t(st["default_prop_states"])
    obj.franka_dof_targets = t(st["franka_dof_targets"])
    obj.franka_default_dof_pos = torch.tensor(cfg["default_dof_pos"], dtype=torch.float32)
    obj.franka_dof_lower_limits, obj.franka_dof_upper_limits = t(syn.FRANKA_DOF_LOWER), t(syn.FRANKA_DOF_UPPER)
    obj.franka_dof_speed_scales = torch.tensor(cfg["dof_speed_scales"], dtype=torch.float32)
    rep = lambda k: torch.tensor(cfg[k], dtype=torch.float32).repeat((n, 1))                      # noqa: E731
    for k in ("franka_local_grasp_pos", "franka_local_grasp_rot", "drawer_local_grasp_pos", "drawer_local_grasp_rot",
              "gripper_forward_axis", "drawer_inward_axis", "gripper_up_axis", "drawer_up_axis"):
        setattr(obj, k, rep(k))
    obj.franka_grasp_pos, obj.drawer_grasp_pos = torch.zeros(n, 3), torch.zeros(n, 3)
    obj.franka_grasp_rot, obj.drawer_grasp_rot = torch.zeros(n, 4), torch.zeros(n, 4)
    obj.global_indices = torch.arange(n * cfg["num_actors"], dtype=torch.int32).view(n, -1)
    return obj


_ATTR = {"rigid_body_state": "_rigid_body_state", "root_state": "_root_state", "dof_state": "_dof_state",
         "pos_control": "_pos_control", "effort_control": "_effort_control", "init_cubeA_state": "_init_cubeA_state",
         "init_cubeB_state": "_init_cubeB_state"}


def _get(obj, task, name):
    return getattr(obj, name if task == "franka_cabinet" else _ATTR.get(name, name))


def replay_draws(task, obj, ids, seed):
    """The uniforms reset_idx consumed, re-drawn from the same seed in the reference's call order; FrankaCubeStack's
    check_valid loop (franka_cube_stack.py:567-585) is followed draw by draw and the ACCEPTED sample kept."""
    k = int(ids.numel())
    cols = FH.draw_cols(task)
    if k == 0:
        return [np.zeros((0, c), np.float32) for c in cols]
    torch.manual_seed(seed)
    if task != "franka_sa_stack":
        return [torch.rand(k, c).numpy() if c > 1 else torch.rand(k).numpy().reshape(k, 1) for c in cols]
    cfg = FH.cfg_of(task)
    ub_xy, ub_yaw = torch.rand(k, 2), torch.rand(k)
    center = torch.tensor(cfg["table_surface_pos"][:2], dtype=torch.float32)
    other = obj._init_cubeB_state[ids, :2]
    min_d = (obj.states["cubeA_size"] + obj.states["cubeB_size"])[ids] * np.sqrt(2) / 2.0 * 2.0
    sampled, acc = torch.zeros(k, 2), torch.zeros(k, 2)
    active = torch.arange(k)
    tries = 0
    for _ in range(100):
        u = torch.rand_like(sampled[active])
        acc[active] = u
        sampled[active] = center + 2.0 * cfg["start_position_noise"] * (u - 0.5)
        tries += 1
        active = torch.nonzero(torch.linalg.norm(sampled - other, dim=-1) < min_d, as_tuple=True)[0]
        if len(active) == 0:
            break
    assert torch.equal(sampled, obj._init_cubeA_state[ids, :2]), "check_valid replay diverged from the reference"
    ua_yaw, uj = torch.rand(k), torch.rand(k, 9)
    replay_draws.tries = tries
    return [ub_xy.numpy(), ub_yaw.numpy().reshape(k, 1), acc.numpy(), ua_yaw.numpy().reshape(k, 1), uj.numpy()]


def generate(task, n, seed, reset_prob, steps):
    st = syn.make_state(task, n, seed=seed, reset_prob=reset_prob)
    L = FH.cfg_of(task)["max_episode_length"]
    st["progress_buf"][1::16] = L - 2
    st["progress_buf"][2::16] = L - 3
    obj = build_cabinet(st) if task == "franka_cabinet" else build_sa(task, st)
    obj.reset_buf, obj.progress_buf = t(st["reset_buf"]), t(st["progress_buf"])
    g = np.random.Generator(np.random.PCG64(seed + 177))
    blob = {"meta_num_envs": n, "meta_steps": steps}
    for k, v in st.items():
        if not k.startswith("cfg_"):
            blob["init/" + k] = v.copy()
    info = []
    for s in range(steps):
        if s > 0:
            syn.advance_physics(task, st, g)
        ids = obj.reset_buf.nonzero(as_tuple=False).flatten().clone()
        mid = {}

        def on_simulate():
            if task == "franka_cabinet":
                mid["franka_dof_targets"] = obj.franka_dof_targets.detach().clone().numpy()
            for name in FH.PHYS[task]:
                _get(obj, task, name).copy_(t(st[name]))
                blob[f"step{s}/phys/{name}"] = st[name].copy()

        obj.gym.on_simulate = on_simulate
        obj.gym.calls.clear()
        obj.extras = {}
        actions = t(st["actions"])
        torch.manual_seed(4200 + s)
        obs_dict, rew, reset, extras = obj.step(actions)           # <- the reference's own code, end to end
        rands = replay_draws(task, obj, ids, 4200 + s)
        for i, r in enumerate(rands):
            blob[f"step{s}/rand{i}"] = np.ascontiguousarray(r, np.float32)
        blob[f"step{s}/actions_clamped"] = torch.clamp(actions, -obj.clip_actions, obj.clip_actions).numpy()
        for name, v in mid.items():
            blob[f"step{s}/mid/{name}"] = v
        for name in FH.OUT[task] + ("reset_buf", "progress_buf"):
            if name != "obs_clamped":
                blob[f"step{s}/out/{name}"] = _get(obj, task, name).detach().clone().numpy()
        blob[f"step{s}/out/obs_clamped"] = obs_dict["obs"].detach().clone().numpy()
        blob[f"step{s}/out/time_outs"] = extras["time_outs"].detach().clone().numpy()
        blob[f"step{s}/out/reset_env_ids"] = ids.to(torch.int32).numpy()
        comps = sorted(k for k in extras if k.startswith("r/") or k in ("mean_cube_height", "target_dist"))
        if comps:
            blob[f"step{s}/out/component_names"] = np.array(comps)
            blob[f"step{s}/out/components"] = np.array([float(extras[c]) for c in comps], np.float32)
        # the sim-call id lists (global actor indices) must be what the env id list implies
        calls = [(name, [a for a in args if isinstance(a, torch.Tensor) and a.dtype == torch.int32]) for name, args in obj.gym.calls]
        if ids.numel() > 0:
            gi = obj.global_indices if task == "franka_cabinet" else obj._global_indices
            dof_ids = [a[0] for name, a in calls if name == "set_dof_state_tensor_indexed"][0]
            want = gi[ids, :2].flatten() if task == "franka_cabinet" else gi[ids, 0].flatten()
            assert torch.equal(dof_ids, want)
        info.append(int(ids.numel()))
    return blob, info


def main():
    plan = [(40, 0, 0.1, 3), (32, 1, 1.0, 2), (56, 2, 0.04, 2)]
    for task in FH.TASKS:
        for n, seed, p, steps in plan:
            blob, info = generate(task, n, seed, p, steps)
            path = os.path.join(HERE, f"{task}_n{n}_s{seed}.npz")
            np.savez_compressed(path, **blob)
            last = steps - 1
            print(f"{os.path.basename(path)}: {os.path.getsize(path) // 1024} KiB, resets/step {info}, done frac "
                  f"{float(blob[f'step{last}/out/reset_buf'].mean()):.2f}, time_outs {int(blob[f'step{last}/out/time_outs'].sum())}"
                  + (f", check_valid tries {getattr(replay_draws, 'tries', 0)}" if task == "franka_sa_stack" else ""))


if __name__ == "__main__":
    main()
