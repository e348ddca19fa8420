"""Class-level oracle harness: drive the UNMODIFIED reference task classes on synthetic state.

Container-only test infrastructure (needs /root/reference).  ``Task.__new__`` + hand-set
attributes + ``FakeGym``, then the reference's own ``VecTask.step`` is
called, which runs pre_physics_step -> (no-op simulate) -> post_physics_step -> the step
epilogue (vec_task.py:362-410).
"""
import warnings

import numpy as np
import torch

import ref_shim

warnings.filterwarnings("ignore", category=SyntaxWarning)


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a).copy())


def _common(obj, cfg, n, num_obs, num_actions, num_states=0, num_agents=1):
    obj.gym = ref_shim.FakeGym()
    obj.sim = object()
    obj.device = "cpu"
    obj.rl_device = "cpu"
    obj.num_environments = n
    obj.num_agents = num_agents
    obj.num_observations = num_obs
    obj.num_states = num_states
    obj.num_actions = num_actions
    obj.control_freq_inv = 1
    obj.clip_obs = cfg.get("clip_obs", float("inf"))
    obj.clip_actions = cfg.get("clip_actions", 1.0)
    obj.dr_randomizations = {}
    obj.force_render = False
    obj.control_steps = 0
    obj.obs_dict = {}
    obj.extras = {}
    obj.viewer = None
    obj.debug_viz = False
    obj.randomize = False
    obj.max_episode_length = cfg["max_episode_length"]
    obj.states_buf = torch.zeros((n, num_states))
    obj.timeout_buf = torch.zeros(n * num_agents, dtype=torch.long)


def build_cartpole(st, cfg):
    mod = ref_shim.ref_module("tasks.cartpole")
    n = st["progress_buf"].shape[0]
    obj = mod.Cartpole.__new__(mod.Cartpole)
    _common(obj, cfg, n, 4, 1)
    obj.reset_dist = cfg["reset_dist"]
    obj.max_push_effort = 400.0
    obj.num_dof = 2
    obj.dof_state = _t(st["dof_state"])
    obj.dof_pos = obj.dof_state.view(n, 2, 2)[..., 0]
    obj.dof_vel = obj.dof_state.view(n, 2, 2)[..., 1]
    obj.obs_buf = _t(st["obs_buf"])
    obj.rew_buf = _t(st["rew_buf"])
    obj.reset_buf = _t(st["reset_buf"])
    obj.progress_buf = _t(st["progress_buf"])
    return obj


def _build_locomotion(cls, st, cfg):
    n = st["progress_buf"].shape[0]
    d = cfg["num_dof"]
    obj = cls.__new__(cls)
    _common(obj, cfg, n, cfg["num_obs"], d)
    for k in ("dof_vel_scale", "contact_force_scale", "power_scale", "heading_weight", "up_weight",
              "actions_cost_scale", "energy_cost_scale", "joints_at_limit_cost_scale", "death_cost",
              "termination_height", "dt"):
        setattr(obj, k, cfg[k])
    obj.up_axis_idx = 2
    obj.num_dof = d
    for k in ("root_states", "initial_root_states", "dof_state", "initial_dof_pos",
              "vec_sensor_tensor", "targets", "inv_start_rot", "basis_vec0", "basis_vec1",
              "potentials", "prev_potentials", "up_vec", "heading_vec", "obs_buf", "rew_buf",
              "reset_buf", "progress_buf", "randomize_buf", "dof_limits_lower", "dof_limits_upper"):
        setattr(obj, k, _t(st[k]))
    obj.dof_pos = obj.dof_state.view(n, d, 2)[..., 0]
    obj.dof_vel = obj.dof_state.view(n, d, 2)[..., 1]
    obj.actions = _t(st["actions"])
    return obj


def build_ant(st, cfg):
    mod = ref_shim.ref_module("tasks.ant")
    obj = _build_locomotion(mod.Ant, st, cfg)
    obj.joint_gears = torch.full((cfg["num_dof"],), 15.0)
    return obj


def build_humanoid(st, cfg):
    mod = ref_shim.ref_module("tasks.humanoid")
    obj = _build_locomotion(mod.Humanoid, st, cfg)
    obj.angular_velocity_scale = cfg["angular_velocity_scale"]
    obj.dof_force_tensor = _t(st["dof_force_tensor"])
    obj.motor_efforts = _t(st["motor_efforts"])
    obj.max_motor_effort = float(obj.motor_efforts.max())
    return obj


BUILDERS = {"cartpole": build_cartpole, "ant": build_ant, "humanoid": build_humanoid}

# buffers the path mutates / produces, per task (captured after every step)
OUTPUTS = {
    "cartpole": ["obs_buf", "rew_buf", "reset_buf", "progress_buf", "dof_state"],
    "ant": ["obs_buf", "rew_buf", "reset_buf", "progress_buf", "randomize_buf", "dof_state",
            "potentials", "prev_potentials", "up_vec", "heading_vec"],
    "humanoid": ["obs_buf", "rew_buf", "reset_buf", "progress_buf", "randomize_buf", "dof_state",
                 "potentials", "prev_potentials", "up_vec", "heading_vec"],
}
# sim-owned tensors refreshed by the synthetic "physics" before every step
SIM_INPUTS = {
    "cartpole": ["dof_state"],
    "ant": ["root_states", "dof_state", "vec_sensor_tensor", "actions"],
    "humanoid": ["root_states", "dof_state", "vec_sensor_tensor", "dof_force_tensor", "actions"],
}


def reset_rand_shapes(task, k, cfg):
    """torch.rand call shapes the reference's reset_idx makes, in order."""
    if task == "cartpole":
        return [(k, 2), (k, 2)]                       # cartpole.py:145-146
    if task in ("ant", "humanoid"):
        return [(k, cfg["num_dof"]), (k, cfg["num_dof"])]   # ant.py:257-258, humanoid.py:258-259
    raise KeyError(task)


def run_reference_step(task, obj, st, cfg, torch_seed):
    """Copy the sim-owned tensors of ``st`` into the reference object, run one reference
    ``VecTask.step`` and return (outputs dict, reset-rand list, recorded sim calls)."""
    n = obj.progress_buf.shape[0]
    for k in SIM_INPUTS[task]:
        if k == "actions":
            continue
        getattr(obj, k).copy_(_t(st[k]))
    reset_ids = obj.reset_buf.nonzero(as_tuple=False).flatten().clone()
    if task == "cartpole":
        actions = torch.zeros(n, 1)
    else:
        actions = _t(st["actions"])
    obj.gym.calls.clear()
    torch.manual_seed(torch_seed)
    obs_dict, rew, reset, extras = obj.step(actions)
    # regenerate the uniforms reset_idx consumed (same seed, same call shapes, same order)
    k = int(reset_ids.numel())
    rands = []
    if k > 0:
        torch.manual_seed(torch_seed)
        rands = [torch.rand(*s).numpy() for s in reset_rand_shapes(task, k, cfg)]
    out = {name: getattr(obj, name).detach().clone().numpy() for name in OUTPUTS[task]}
    out["obs_clamped"] = obs_dict["obs"].detach().clone().numpy()
    out["time_outs"] = extras["time_outs"].detach().clone().numpy()
    out["reset_env_ids"] = reset_ids.to(torch.int32).numpy()
    calls = [(name, [a.numpy() if isinstance(a, torch.Tensor) else a for a in args])
             for name, args in obj.gym.calls]
    return out, rands, calls
