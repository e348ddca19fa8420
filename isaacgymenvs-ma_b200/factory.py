"""Build task mirrors on synthetic simulator state (the stand-in for isaacgymenvs.make(), whose
hydra/isaacgym machinery is outside the hot path; signature kept close: seed, task, num_envs,
sim_device, rl_device)."""
import numpy as np
import torch

from . import synthetic
from .tasks import SyntheticSim, task_map

_SIM_TENSORS = {
    "Cartpole": ["dof_state"],
    "Ant": ["root_states", "dof_state", "vec_sensor_tensor"],
    "Humanoid": ["root_states", "dof_state", "vec_sensor_tensor", "dof_force_tensor"],
    "ShadowHand": ["dof_state", "root_state_tensor", "rigid_body_states", "vec_sensor_tensor", "dof_force_tensor"],
}
_SIM_TENSORS["Anymal"] = ["root_states", "dof_state", "contact_forces", "torques"]
_MA_TASKS = {"FrankaReachMA": "reach", "FrankaCollectMA": "collect", "FrankaCombineMA": "combine", "FrankaPPMA": "pp"}
for _t in _MA_TASKS:
    _SIM_TENSORS[_t] = ["_rigid_body_state", "_root_state", "_dof_state", "_contact_forces"]
_CFGS = {"Cartpole": synthetic.CARTPOLE_CFG, "Ant": synthetic.ANT_CFG, "Humanoid": synthetic.HUMANOID_CFG,
         "ShadowHand": synthetic.SHADOW_HAND_CFG, "Anymal": synthetic.ANYMAL_CFG}
_SYN_NAME = {"Cartpole": "cartpole", "Ant": "ant", "Humanoid": "humanoid", "ShadowHand": "shadow_hand", "Anymal": "anymal"}
for _t, _v in _MA_TASKS.items():
    _SYN_NAME[_t] = "franka_" + _v


def task_cfg(task):
    if task in _MA_TASKS:
        return synthetic.franka_ma_cfg(_MA_TASKS[task])
    return dict(_CFGS[task])


def sim_tensor_names(task):
    return list(_SIM_TENSORS[task])


def from_state(task, st, sim_device, rl_device=None, **kw):
    """Construct the mirror of `task` around an existing synthetic state dict (numpy arrays)."""
    dev = torch.device(sim_device)
    cfg = task_cfg(task)
    if task in _MA_TASKS:
        cfg = synthetic.franka_ma_cfg(_MA_TASKS[task], int(st["cfg_num_agents"]), int(st["cfg_num_targets"]))
        cfg["num_envs"] = int(st["progress_buf"].shape[0]) // cfg["num_agents"]
    else:
        cfg["num_envs"] = int(st["progress_buf"].shape[0])
    for k in list(st):                      # per-state config overrides travel with the state dict
        if k.startswith("cfg_"):
            v = st[k]
            if task not in _MA_TASKS:
                cfg[k[4:]] = v.item() if v.shape == () and v.dtype.kind in "iuf" else str(v)
    tensors = {k: torch.from_numpy(np.ascontiguousarray(st[k])).to(dev) for k in _SIM_TENSORS[task]}
    sim = SyntheticSim(tensors)
    obj = task_map[task](cfg, sim, dev, rl_device=rl_device, **kw)
    if hasattr(obj, "adopt_state"):
        obj.adopt_state(st)
    else:
        for k in ("reset_buf", "progress_buf"):
            getattr(obj, k).copy_(torch.from_numpy(st[k]).to(dev))
    return obj


def make(seed, task, num_envs, sim_device, rl_device=None, reset_prob=0.02, **kw):
    """isaacgymenvs.make()-shaped entry point (isaacgymenvs/__init__.py:14-26) on synthetic state."""
    st = synthetic.make_state(_SYN_NAME[task], num_envs, seed=seed, reset_prob=reset_prob)
    kw.setdefault("seed", seed)
    return from_state(task, st, sim_device, rl_device=rl_device, **kw)


def load_sim_inputs(obj, task, arrays):
    """Overwrite the simulator-owned tensors (and actions) with new post-simulate values."""
    for k, v in arrays.items():
        t = obj.sim.tensors.get(k)
        if t is None:
            t = getattr(obj, k)
        t.copy_(torch.from_numpy(np.ascontiguousarray(v)).to(t.device).view_as(t))
