"""Golden fixtures for Anymal: the UNMODIFIED reference ``Anymal`` class
(/root/reference/isaacgymenvs/tasks/anymal.py) stepped through ``VecTask.step`` on synthetic state, built with
``Task.__new__`` + ``FakeGym`` exactly like tests/golden/gen_golden.py does for the other tasks.

Container-only (needs /root/reference):  python tests/golden/gen_golden_anymal.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402
import ref_shim  # noqa: E402

from gen_golden_anymal_state import CFG, DT, advance, make_state  # noqa: E402


def build(st):
    mod = ref_shim.ref_module("tasks.anymal")
    n = st["reset_buf"].shape[0]
    obj = mod.Anymal.__new__(mod.Anymal)
    ref_harness._common(obj, CFG, n, 48, 12)
    t = ref_harness._t
    for k in ("lin_vel_scale", "ang_vel_scale", "dof_pos_scale", "dof_vel_scale", "action_scale",
              "command_x_range", "command_y_range", "command_yaw_range", "base_index"):
        setattr(obj, k, CFG[k])
    obj.rew_scales = {"lin_vel_xy": CFG["rew_lin_vel_xy"] * DT, "ang_vel_z": CFG["rew_ang_vel_z"] * DT,
                      "torque": CFG["rew_torque"] * DT}                      # anymal.py:99-100
    obj.num_dof = 12
    obj.root_states, obj.dof_state = t(st["root_states"]), t(st["dof_state"])
    obj.dof_pos, obj.dof_vel = obj.dof_state.view(n, 12, 2)[..., 0], obj.dof_state.view(n, 12, 2)[..., 1]
    obj.contact_forces, obj.torques = t(st["contact_forces"]), t(st["torques"])
    obj.commands = t(st["commands"])
    obj.commands_x, obj.commands_y, obj.commands_yaw = (obj.commands.view(n, 3)[..., i] for i in range(3))
    obj.default_dof_pos = torch.tensor(CFG["default_dof_pos"]).repeat(n, 1)
    obj.initial_root_states = obj.root_states.clone()
    obj.gravity_vec = torch.tensor([0.0, 0.0, -1.0]).repeat(n, 1)
    obj.knee_indices = torch.tensor(CFG["knee_indices"], dtype=torch.long)
    obj.actions = torch.zeros(n, 12)
    obj.obs_buf, obj.rew_buf = torch.zeros(n, 48), torch.zeros(n)
    obj.reset_buf, obj.progress_buf = t(st["reset_buf"]), t(st["progress_buf"])
    obj.randomize_buf = torch.zeros(n, dtype=torch.long)
    return obj


def generate(n, seed, reset_prob, steps):
    st = make_state(n, seed, reset_prob)
    obj = build(st)
    g = np.random.Generator(np.random.PCG64(seed + 99))
    blob = {"meta_num_envs": n, "meta_steps": steps, "init/reset_buf": st["reset_buf"], "init/progress_buf": st["progress_buf"],
            "init/commands": st["commands"]}
    for k, v in CFG.items():
        blob["cfg_" + k] = np.asarray(v)
    for s in range(steps):
        if s > 0:
            st["dof_state"] = obj.dof_state.numpy().copy()
            advance(st, g)
        for k in ("root_states", "dof_state", "contact_forces", "torques"):
            getattr(obj, k).copy_(ref_harness._t(st[k]))
            blob[f"step{s}/in/{k}"] = st[k].copy()
        blob[f"step{s}/in/actions"] = st["actions"].copy()
        ids = obj.reset_buf.nonzero(as_tuple=False).flatten().clone()
        k = int(ids.numel())
        obj.gym.calls.clear()
        torch.manual_seed(500 + s)
        obs_dict, rew, reset, extras = obj.step(ref_harness._t(st["actions"]))   # <- the reference's own code
        if k > 0:
            torch.manual_seed(500 + s)
            for i, shape in enumerate([(k, 12), (k, 12), (k, 1), (k, 1), (k, 1)]):   # anymal.py:282-302
                blob[f"step{s}/rand{i}"] = torch.rand(*shape).numpy()
        for name in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "dof_state", "commands", "actions"):
            blob[f"step{s}/out/{name}"] = getattr(obj, name).detach().clone().numpy()
        blob[f"step{s}/out/obs_clamped"] = obs_dict["obs"].detach().clone().numpy()
        blob[f"step{s}/out/time_outs"] = extras["time_outs"].detach().clone().numpy()
        blob[f"step{s}/out/reset_env_ids"] = ids.to(torch.int32).numpy()
        idx = [a for name, args in obj.gym.calls for a in args if isinstance(a, torch.Tensor) and a.dtype == torch.int32]
        if idx:
            assert all(np.array_equal(i.numpy(), ids.to(torch.int32).numpy()) for i in idx)
    return blob


def main():
    for n, seed, p, steps in [(96, 0, 0.1, 3), (64, 1, 1.0, 2), (300, 2, 0.02, 2)]:
        blob = generate(n, seed, p, steps)
        path = os.path.join(HERE, f"anymal_n{n}_s{seed}.npz")
        np.savez_compressed(path, **blob)
        last = steps - 1
        print(f"{os.path.basename(path)}: {os.path.getsize(path) // 1024} KiB, resets/step "
              f"{[int(blob[f'step{s}/out/reset_env_ids'].size) for s in range(steps)]}, "
              f"done frac {float(blob[f'step{last}/out/reset_buf'].mean()):.2f}, "
              f"rew>0 frac {float((blob[f'step{last}/out/rew_buf'] > 0).mean()):.2f}")


if __name__ == "__main__":
    main()
