"""Golden fixtures for BallBalance / Quadcopter / Ingenuity: the UNMODIFIED reference classes
(/root/reference/isaacgymenvs/tasks/{ball_balance,quadcopter,ingenuity}.py) stepped through ``VecTask.step`` on
synthetic state, built with ``Task.__new__`` + ``FakeGym`` like tests/golden/gen_golden.py does for the other tasks.

These tasks reset inside ``pre_physics_step``, so ``gym.simulate`` is hooked: the hook snapshots what pre-physics wrote
("mid/"), then loads the next synthetic simulator state ("phys/") before ``post_physics_step`` runs.

Container-only (needs /root/reference):  python tests/golden/gen_golden_small.py
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
import small_harness as SH  # noqa: E402

syn = H.pkg.synthetic
t = ref_harness._t


class HookGym(ref_shim.FakeGym):
    def __init__(self):
        super().__init__()
        self.on_simulate = None

    def simulate(self, sim):
        if self.on_simulate:
            self.on_simulate()


def _base(cls, cfg, n, num_obs, num_actions):
    obj = cls.__new__(cls)
    ref_harness._common(obj, cfg, n, num_obs, num_actions)
    obj.gym = HookGym()
    obj.dt = cfg["dt"]
    obj.obs_buf, obj.rew_buf = torch.zeros(n, num_obs), torch.zeros(n)
    obj.randomize_buf = torch.zeros(n, dtype=torch.long)
    return obj


def build_ball_balance(st):
    cfg, n = syn.BALL_BALANCE_CFG, st["reset_buf"].shape[0]
    obj = _base(ref_shim.ref_module("tasks.ball_balance").BallBalance, cfg, n, 24, 3)
    obj.action_speed_scale = cfg["action_speed_scale"]
    obj.ball_radius = cfg["ball_radius"]
    root, dof, sens = t(st["root_states"]), t(st["dof_states"]), t(st["sensor_tensor"])
    obj.root_tensor, obj.dof_state_tensor, obj.sensor_tensor = root, dof, sens     # ball_balance.py:87-112
    obj.root_states = root
    obj.tray_positions, obj.ball_positions = root[..., 0, 0:3], root[..., 1, 0:3]
    obj.ball_orientations, obj.ball_linvels, obj.ball_angvels = root[..., 1, 3:7], root[..., 1, 7:10], root[..., 1, 10:13]
    obj.dof_states, obj.dof_positions, obj.dof_velocities = dof, dof[..., 0], dof[..., 1]
    obj.sensor_forces, obj.sensor_torques = sens[..., 0:3], sens[..., 3:6]
    obj.initial_dof_states, obj.initial_root_states = t(st["initial_dof_states"]), t(st["initial_root_states"])
    obj.dof_position_targets = t(st["dof_position_targets"])
    obj.all_actor_indices = torch.arange(2 * n, dtype=torch.int32).view(n, 2)
    obj.all_bbot_indices = 2 * torch.arange(n, dtype=torch.int32)
    obj.bbot_dof_lower_limits = torch.tensor(cfg["dof_lower"], dtype=torch.float32)   # to_torch: fp32
    obj.bbot_dof_upper_limits = torch.tensor(cfg["dof_upper"], dtype=torch.float32)
    return obj


def build_quadcopter(st):
    cfg, n = syn.QUADCOPTER_CFG, st["reset_buf"].shape[0]
    obj = _base(ref_shim.ref_module("tasks.quadcopter").Quadcopter, cfg, n, 21, 12)
    root, dof = t(st["root_states"]), t(st["dof_states"])
    obj.root_tensor, obj.dof_state_tensor = root, dof                               # quadcopter.py:66-84
    obj.root_states, obj.root_positions, obj.root_quats = root, root[..., 0:3], root[..., 3:7]
    obj.root_linvels, obj.root_angvels = root[..., 7:10], root[..., 10:13]
    obj.dof_states, obj.dof_positions, obj.dof_velocities = dof, dof[..., 0], dof[..., 1]
    obj.initial_root_states, obj.initial_dof_states = t(st["initial_root_states"]), t(st["initial_dof_states"])
    obj.thrust_lower_limits = torch.zeros(4)
    obj.thrust_upper_limits = cfg["max_thrust"] * torch.ones(4)
    obj.dof_lower_limits = torch.tensor(cfg["dof_lower"], dtype=torch.float32)
    obj.dof_upper_limits = torch.tensor(cfg["dof_upper"], dtype=torch.float32)
    obj.dof_position_targets, obj.thrusts, obj.forces = t(st["dof_position_targets"]), t(st["thrusts"]), t(st["forces"])
    obj.all_actor_indices = torch.arange(n, dtype=torch.int32)
    return obj


def build_ingenuity(st):
    cfg, n = syn.INGENUITY_CFG, st["reset_buf"].shape[0]
    obj = _base(ref_shim.ref_module("tasks.ingenuity").Ingenuity, cfg, n, 13, 6)
    vec_root, dof = t(st["root_tensor"]), t(st["dof_states"])
    obj.root_tensor, obj.dof_state_tensor = vec_root, dof                           # ingenuity.py:63-92
    obj.root_states = vec_root[:, 0, :]
    obj.root_positions, obj.root_quats = obj.root_states[:, 0:3], obj.root_states[:, 3:7]
    obj.root_linvels, obj.root_angvels = obj.root_states[:, 7:10], obj.root_states[:, 10:13]
    obj.target_root_positions = t(st["target_root_positions"])
    obj.marker_states = vec_root[:, 1, :]
    obj.marker_positions = obj.marker_states[:, 0:3]
    obj.dof_states, obj.dof_positions, obj.dof_velocities = dof, dof[..., 0], dof[..., 1]
    obj.initial_root_states = t(st["initial_root_states"])
    obj.thrust_lower_limit, obj.thrust_upper_limit = 0, int(cfg["thrust_upper_limit"])
    obj.thrust_lateral_component = cfg["thrust_lateral_component"]
    obj.thrusts, obj.forces = t(st["thrusts"]), t(st["forces"])
    obj.all_actor_indices = torch.arange(n * 2, dtype=torch.int32).reshape(n, 2)
    return obj


BUILD = {"ball_balance": build_ball_balance, "quadcopter": build_quadcopter, "ingenuity": build_ingenuity}


def generate(task, n, seed, reset_prob, steps):
    st = syn.make_state(task, n, seed=seed, reset_prob=reset_prob)
    st["progress_buf"][1::16] = SH.CFG[task]["max_episode_length"] - 2      # time-outs in the first step
    st["progress_buf"][2::16] = SH.CFG[task]["max_episode_length"] - 3      # ... and in the second
    obj = BUILD[task](st)
    obj.reset_buf, obj.progress_buf = t(st["reset_buf"]), t(st["progress_buf"])
    g = np.random.Generator(np.random.PCG64(seed + 77))
    blob = {"meta_num_envs": n, "meta_steps": steps}
    for k, v in st.items():
        blob["init/" + k] = v.copy()
    for s in range(steps):
        ids = obj.reset_buf.nonzero(as_tuple=False).flatten().clone()
        k = int(ids.numel())
        kt = int((obj.progress_buf % 500 == 0).sum()) if task == "ingenuity" else 0
        mid = {}

        def on_simulate():
            for name in SH.PRE_OUT[task] + ("reset_buf", "progress_buf"):
                mid[name] = getattr(obj, name).detach().clone().numpy()
            syn.advance_physics(task, st, g)                       # the simulator's new state
            for name in SH.PHYS[task]:
                getattr(obj, name).copy_(t(st[name]))
                blob[f"step{s}/phys/{name}"] = st[name].copy()

        obj.gym.on_simulate = on_simulate
        obj.gym.calls.clear()
        actions = t(st["actions"])
        torch.manual_seed(900 + s)
        obs_dict, rew, reset, extras = obj.step(actions)           # <- the reference's own code, end to end
        torch.manual_seed(900 + s)
        for i, shape in enumerate(SH.draw_shapes(task, k, kt)):    # the same stream, in the reference's call order
            if task == "ingenuity" and i == 1:
                shape = (shape[0],)                                # torch.rand(num_sets) ingenuity.py:287
            if task == "ingenuity" and i == 3:
                shape = (shape[0],)
            blob[f"step{s}/rand{i}"] = (torch.rand(*shape).numpy().reshape(shape[0], -1) if shape[0] > 0
                                        else np.zeros((0, SH.small.DRAW_COLS[task][i]), np.float32))
        blob[f"step{s}/actions"] = st["actions"].copy()
        blob[f"step{s}/actions_clamped"] = torch.clamp(actions, -obj.clip_actions, obj.clip_actions).numpy()
        for name, v in mid.items():
            blob[f"step{s}/mid/{name}"] = v
        blob[f"step{s}/mid/reset_env_ids"] = ids.to(torch.int32).numpy()
        idx = [(name, a) for name, args in obj.gym.calls for a in args if isinstance(a, torch.Tensor) and a.dtype == torch.int32]
        if task == "ingenuity":
            # :341 the sorted unique() of the target-marker and copter actor rows
            root_ids = [a.numpy() for name, a in idx if name == "set_actor_root_state_tensor_indexed"]
            blob[f"step{s}/mid/actor_ids"] = root_ids[0] if root_ids else np.zeros(0, np.int32)
        elif task == "quadcopter" and k > 0:
            assert all(np.array_equal(a.numpy(), ids.to(torch.int32).numpy()) for _, a in idx)
        for name in ("obs_buf", "rew_buf", "reset_buf", "progress_buf"):
            blob[f"step{s}/out/{name}"] = getattr(obj, name).detach().clone().numpy()
        blob[f"step{s}/out/obs_clamped"] = obs_dict["obs"].detach().clone().numpy()
        blob[f"step{s}/out/time_outs"] = extras["time_outs"].detach().clone().numpy()
        # next step: task-owned state carries over inside obj; the synthetic dict follows what the reference holds
        for name in SH.PHYS[task]:
            st[name] = getattr(obj, name).detach().numpy().copy()
        st["actions"] = g.uniform(-1.5, 1.5, st["actions"].shape).astype(np.float32)
    return blob


def main():
    plan = [(64, 0, 0.1, 3), (48, 1, 1.0, 2), (96, 2, 0.02, 3)]
    for task in SH.TASKS:
        for n, seed, p, steps in plan:
            blob = generate(task, n, seed, p, steps)
            path = os.path.join(HERE, f"{task}_n{n}_s{seed}.npz")
            np.savez_compressed(path, **blob)
            print(f"{os.path.basename(path)}: {os.path.getsize(path) // 1024} KiB, resets/step "
                  f"{[int(blob[f'step{s}/mid/reset_env_ids'].size) for s in range(steps)]}, "
                  f"done frac {float(blob[f'step{steps - 1}/out/reset_buf'].mean()):.2f}, "
                  f"time_outs {int(blob[f'step{steps - 1}/out/time_outs'].sum())}")


if __name__ == "__main__":
    main()
