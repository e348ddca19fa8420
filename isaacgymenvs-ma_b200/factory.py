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
_SIM_TENSORS["AllegroHand"] = list(_SIM_TENSORS["ShadowHand"])
_SIM_TENSORS["Anymal"] = ["root_states", "dof_state", "contact_forces", "torques"]
_SIM_TENSORS.update({"BallBalance": ["root_states", "dof_states", "sensor_tensor"], "Quadcopter": ["root_states", "dof_states"],
                     "Ingenuity": ["root_tensor", "dof_states"]})
_SA_TASKS = {"FrankaReach": "reach", "FrankaCubeStack2": "stack2", "FrankaCubeStack": "stack"}
for _t in _SA_TASKS:
    _SIM_TENSORS[_t] = ["rigid_body_state", "root_state", "dof_state"]
_SIM_TENSORS["FrankaCabinet"] = ["rigid_body_states", "root_state_tensor", "dof_state"]
_MA_TASKS = {"FrankaReachMA": "reach", "FrankaCollectMA": "collect", "FrankaCombineMA": "combine", "FrankaPPMA": "pp"}
for _t in _MA_TASKS:
    _SIM_TENSORS[_t] = ["_rigid_body_state", "_root_state", "_dof_state", "_contact_forces"]
_CFGS = {"Cartpole": synthetic.CARTPOLE_CFG, "Ant": synthetic.ANT_CFG, "Humanoid": synthetic.HUMANOID_CFG,
         "ShadowHand": synthetic.SHADOW_HAND_CFG, "Anymal": synthetic.ANYMAL_CFG,
         "BallBalance": synthetic.BALL_BALANCE_CFG, "Quadcopter": synthetic.QUADCOPTER_CFG,
         "Ingenuity": synthetic.INGENUITY_CFG}
_SYN_NAME = {"Cartpole": "cartpole", "Ant": "ant", "Humanoid": "humanoid", "ShadowHand": "shadow_hand", "Anymal": "anymal",
             "BallBalance": "ball_balance", "Quadcopter": "quadcopter", "Ingenuity": "ingenuity"}
for _t, _v in _MA_TASKS.items():
    _SYN_NAME[_t] = "franka_" + _v
for _t, _v in _SA_TASKS.items():
    _SYN_NAME[_t] = "franka_sa_" + _v
_SYN_NAME["FrankaCabinet"] = "franka_cabinet"
_CFGS["FrankaCabinet"] = synthetic.FRANKA_CABINET_CFG
_SYN_NAME["AllegroHand"] = "allegro_hand"
_CFGS["AllegroHand"] = synthetic.ALLEGRO_HAND_CFG


def task_cfg(task):
    if task in _MA_TASKS:
        return synthetic.franka_ma_cfg(_MA_TASKS[task])
    if task in _SA_TASKS:
        return synthetic.franka_sa_cfg(_SA_TASKS[task])
    return dict(_CFGS[task])


def sim_tensor_names(task):
    return list(_SIM_TENSORS[task])


def from_state(task, st, sim_device, rl_device=None, dr_randomizations=None, **kw):
    """Construct the mirror of `task` around an existing synthetic state dict (numpy arrays).
    ``dr_randomizations``: the reference's ``task.randomization_params`` dict; only its ``observations`` / ``actions``
    noise entries (and ``frequency``) are part of the fused path (tasks/base.py)."""
    dev = torch.device(sim_device)
    cfg = task_cfg(task)
    if dr_randomizations:
        cfg["dr_randomizations"] = dict(dr_randomizations)
    if task in _MA_TASKS:
        cfg = synthetic.franka_ma_cfg(_MA_TASKS[task], int(st["cfg_num_agents"]), int(st["cfg_num_targets"]))
        cfg["num_envs"] = int(st["progress_buf"].shape[0]) // cfg["num_agents"]
    else:
        cfg["num_envs"] = int(st["progress_buf"].shape[0])
    for k in list(st):                      # per-state config overrides travel with the state dict
        if k.startswith("cfg_"):
            v = st[k]
            if task not in _MA_TASKS and task not in _SA_TASKS:
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


def make(seed, task, num_envs, sim_device, rl_device=None, graphics_device_id=-1, headless=False, multi_gpu=False,
         virtual_screen_capture=False, force_render=True, cfg=None, reset_prob=0.02, **kw):
    """``isaacgymenvs.make()`` (isaacgymenvs/__init__.py:14-26), same positional and keyword arguments, on synthetic
    simulator state.  ``graphics_device_id``, ``headless``, ``virtual_screen_capture`` and ``force_render`` configure
    the viewer / camera sensors, which do not exist here (no graphics stack, no PhysX): accepted and ignored.
    ``multi_gpu``: as in utils/rlgames_utils.py:89-107, rank / device / seed come from the torchrun environment
    (LOCAL_RANK -> cuda:<local_rank>, seed + rank); ``num_envs`` is per rank like the reference's.
    ``cfg``: the hydra config of the reference; only ``cfg["task"]["env"]`` scalars that the mirrors know
    (``numEnvs``, ``clipObservations``, ``clipActions``, ``episodeLength`` / ``maxEpisodeLength``) are honoured."""
    del graphics_device_id, headless, virtual_screen_capture, force_render
    if task not in task_map:
        raise ValueError(f"unknown task {task!r}; the fused path covers {sorted(task_map)}")
    if multi_gpu:
        from . import distributed
        rank, local_rank, _ = distributed.env_info()
        seed = distributed.rank_seed(seed, rank)
        sim_device = f"cuda:{local_rank}"
        rl_device = f"cuda:{local_rank}"
    overrides = {}
    if cfg is not None:
        task_cfg_in = cfg.get("task", cfg) or {}
        if task_cfg_in.get("task", {}).get("randomize") and "dr_randomizations" not in kw:   # ant.py:48-49
            kw["dr_randomizations"] = task_cfg_in["task"].get("randomization_params", {})
        env_cfg = dict((cfg.get("task", cfg) or {}).get("env", {}))
        num_envs = int(env_cfg.get("numEnvs", num_envs))
        for src, dst in (("clipObservations", "clip_obs"), ("clipActions", "clip_actions"),
                         ("episodeLength", "max_episode_length"), ("maxEpisodeLength", "max_episode_length")):
            if src in env_cfg:
                overrides["cfg_" + dst] = np.asarray(env_cfg[src])
    st = synthetic.make_state(_SYN_NAME[task], num_envs, seed=seed, reset_prob=reset_prob)
    if task not in _MA_TASKS:
        st.update(overrides)
    kw.setdefault("seed", seed)
    return from_state(task, st, sim_device, rl_device=rl_device, **kw)


def load_sim_inputs(obj, task, arrays):
    """Overwrite the simulator-owned tensors (and actions) with new post-simulate values."""
    for k, v in arrays.items():
        t = obj.sim.tensors.get(k)
        if t is None:
            t = getattr(obj, k)
        t.copy_(torch.from_numpy(np.ascontiguousarray(v)).to(t.device).view_as(t))
