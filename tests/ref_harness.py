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


def build_shadow_hand(st, cfg):
    allegro = cfg.get("hand") == "allegro"
    mod = ref_shim.ref_module("tasks.allegro_hand" if allegro else "tasks.shadow_hand")
    cls = mod.AllegroHand if allegro else mod.ShadowHand
    n = st["progress_buf"].shape[0]
    d = cfg["num_hand_dofs"]
    obs_type = str(st["cfg_obs_type"])
    asym = bool(int(st["cfg_asymmetric_obs"]))
    obj = cls.__new__(cls)
    _common(obj, cfg, n, st["obs_buf"].shape[1], cfg["num_actions"], num_states=(88 if allegro else 211) if asym else 0)
    obj.obs_type, obj.asymmetric_obs = obs_type, asym
    obj.object_type = "block"
    obj.up_axis_idx = 2
    obj.num_shadow_hand_dofs = d
    obj.num_fingertips = cfg["num_fingertips"]
    obj.dt = cfg["dt"]
    for k in ("dist_reward_scale", "rot_reward_scale", "rot_eps", "action_penalty_scale", "reach_goal_bonus",
              "fall_dist", "fall_penalty", "success_tolerance", "vel_obs_scale", "force_torque_obs_scale",
              "reset_position_noise", "reset_dof_pos_noise", "reset_dof_vel_noise", "use_relative_control",
              "act_moving_average"):
        setattr(obj, k, cfg[k])
    obj.shadow_hand_dof_speed_scale = cfg["dof_speed_scale"]
    obj.max_consecutive_successes = int(st["cfg_max_consecutive_successes"])
    obj.av_factor = torch.tensor(cfg["av_factor"], dtype=torch.float)       # shadow_hand.py:196
    obj.force_scale = 0.0
    obj.print_success_stat = False
    obj.force_prob_range = torch.tensor(cfg["force_prob_range"], dtype=torch.float)
    obj.goal_displacement_tensor = torch.tensor(cfg["goal_displacement"], dtype=torch.float)
    for k in ("dof_state", "root_state_tensor", "vec_sensor_tensor", "dof_force_tensor", "goal_states",
              "goal_init_state", "object_init_state", "hand_indices", "object_indices", "goal_object_indices",
              "fingertip_handles", "actuated_dof_indices", "shadow_hand_dof_lower_limits",
              "shadow_hand_dof_upper_limits", "shadow_hand_dof_default_pos", "shadow_hand_dof_default_vel",
              "progress_buf", "reset_buf", "reset_goal_buf", "randomize_buf", "successes", "consecutive_successes",
              "prev_targets", "cur_targets", "rb_forces", "random_force_prob", "obs_buf", "states_buf", "rew_buf"):
        setattr(obj, k, _t(st[k]))
    obj.rigid_body_states = _t(st["rigid_body_states"])                      # (N, B, 13) view, shadow_hand.py:173
    obj.shadow_hand_default_dof_pos = torch.zeros(d)                         # shadow_hand.py:166
    obj.shadow_hand_dof_state = obj.dof_state.view(n, -1, 2)[:, :d]
    obj.shadow_hand_dof_pos = obj.shadow_hand_dof_state[..., 0]
    obj.shadow_hand_dof_vel = obj.shadow_hand_dof_state[..., 1]
    obj.x_unit_tensor = torch.tensor([1.0, 0, 0]).repeat((n, 1))
    obj.y_unit_tensor = torch.tensor([0, 1.0, 0]).repeat((n, 1))
    obj.z_unit_tensor = torch.tensor([0, 0, 1.0]).repeat((n, 1))
    obj.actions = _t(st["actions"])
    return obj


_MA_MODULE = {"reach": ("tasks.franka_reach_MA", "FrankaReachMA"), "collect": ("tasks.franka_collect_MA", "FrankaCollectMA"),
              "combine": ("tasks.franka_combine_MA", "FrankaCombineMA"), "pp": ("tasks.franka_PPMA", "FrankaPPMA")}


def build_franka_ma(st, cfg):
    """cfg = synthetic.franka_ma_cfg(variant, A, T).  pre_physics_step is replaced by `self.actions = actions`
    (its OSC torque controller needs PhysX Jacobians / mass matrices and is out of the hot path)."""
    v = cfg["variant"]
    modname, clsname = _MA_MODULE[v]
    cls = getattr(ref_shim.ref_module(modname), clsname)
    a, t = cfg["num_agents"], cfg["num_targets"]
    n = st["progress_buf"].shape[0] // a
    obj = cls.__new__(cls)
    _common(obj, cfg, n, cfg["num_obs"], cfg["num_actions"], num_agents=a)
    obj.num_targets = t
    obj.num_dofs = cfg["num_dofs"]
    obj.states, obj.handles = {}, {}
    obj.cubeA_size, obj.dest_size = cfg["cubeA_size"], cfg["dest_size"]
    obj.start_position_noise, obj.start_rotation_noise = cfg["start_position_noise"], 0.785
    obj.franka_dof_noise = cfg["franka_dof_noise"]
    obj._table_surface_pos = np.array(cfg["table_surface_pos"])
    obj.reward_settings = {"r_dist_scale": 0.1, "r_lift_scale": 1.5, "r_align_scale": 2.0, "r_stack_scale": 16.0,
                           "table_height": obj._table_surface_pos[2]}
    obj.wall_pos, obj.wall_height = [0.0, cfg["wall_pos_y"], 1.175], cfg["wall_height"]
    from ige_loader import load_package
    syn = load_package().synthetic
    FRANKA_DEFAULT_DOF_POS, FRANKA_DOF_LOWER, FRANKA_DOF_UPPER = (syn.FRANKA_DEFAULT_DOF_POS, syn.FRANKA_DOF_LOWER,
                                                                  syn.FRANKA_DOF_UPPER)
    obj.franka_default_dof_pos = _t(FRANKA_DEFAULT_DOF_POS)
    obj.franka_dof_lower_limits, obj.franka_dof_upper_limits = _t(FRANKA_DOF_LOWER), _t(FRANKA_DOF_UPPER)
    for k in ("_rigid_body_state", "_root_state", "_dof_state", "_contact_forces", "_pos_control", "_effort_control",
              "_init_cubeA_state", "obs_buf", "rew_buf", "reset_buf", "progress_buf"):
        setattr(obj, k, _t(st[k]))
    obj._q, obj._qd = obj._dof_state[..., 0], obj._dof_state[..., 1]
    rb = obj._rigid_body_state

    def strided(tensor, start, step, count):
        return tensor[:, start:start + step * count:step, :]
    obj._eef_state = strided(rb, cfg["eef_body_start"], cfg["eef_body_step"], a)
    obj._base_state = strided(rb, cfg["base_body_start"], cfg["base_body_step"], a)
    obj._eef_lf_state = obj._eef_state.clone()
    obj._eef_rf_state = obj._eef_state.clone()
    obj._hands_contact_forces = strided(obj._contact_forces, cfg["hand_body_start"], cfg["hand_body_step"], a)
    cs = cfg["cube_actor_start"]
    obj._cubeA_ids = list(range(cs, cs + t))
    obj._cubeA_state = strided(obj._root_state, cs, cfg["cube_actor_step"], t)
    obj.frankas = [list(range(cfg["franka_actor_start"], cfg["franka_actor_start"] + a))]
    obj._global_indices = torch.arange(n * cfg["num_actors"], dtype=torch.int32).view(n, -1)
    obj.states["cubeA_size"] = torch.ones(n, 1) * cfg["cubeA_size"]
    if v == "pp":
        ds = cfg["dest_actor_start"]
        obj._dest_ids = list(range(ds, ds + t))
        obj._dest_state = strided(obj._root_state, ds, cfg["dest_actor_step"], t)
    if v == "combine":
        obj._stack_base_id = cfg["stack_base_actor"]
        obj._stack_base_state = obj._root_state[:, obj._stack_base_id, :]
        obj.agent_idx = torch.zeros(n * a, 1, dtype=torch.long)             # franka_combine_MA.py:526-527
        obj.agent_idx[a - 1::a, :] = 1
        obj.states["agent_idx"] = obj.agent_idx
    obj.actions = _t(st["actions"])
    obj.pre_physics_step = lambda act: setattr(obj, "actions", act.clone())
    return obj


BUILDERS = {"cartpole": build_cartpole, "ant": build_ant, "humanoid": build_humanoid,
            "shadow_hand": build_shadow_hand, "allegro_hand": build_shadow_hand}
for _v in ("reach", "collect", "combine", "pp"):
    BUILDERS["franka_" + _v] = build_franka_ma


def franka_rand_shapes(variant, k, a, t):
    """torch.rand call shapes of reset_idx, in order (franka_reach_MA.py:616-704, franka_PPMA.py:776-895,
    franka_combine_MA.py:793-908)."""
    shapes = [(k, t), (k, t, 2)]                 # _reset_init_cube_state: z then xy
    if variant == "pp":
        shapes += [(k, t), (k, t, 2)]            # _reset_dest_cube_state
    if variant == "combine":
        shapes += [(k,), (k, 2)]                 # _reset_stack_base_cube_state
    shapes.append((k, 9))                        # franka dof noise
    return shapes

# buffers the path mutates / produces, per task (captured after every step)
OUTPUTS = {
    "cartpole": ["obs_buf", "rew_buf", "reset_buf", "progress_buf", "dof_state"],
    "ant": ["obs_buf", "rew_buf", "reset_buf", "progress_buf", "randomize_buf", "dof_state",
            "potentials", "prev_potentials", "up_vec", "heading_vec"],
    "humanoid": ["obs_buf", "rew_buf", "reset_buf", "progress_buf", "randomize_buf", "dof_state",
                 "potentials", "prev_potentials", "up_vec", "heading_vec"],
}
OUTPUTS["shadow_hand"] = ["obs_buf", "states_buf", "rew_buf", "reset_buf", "reset_goal_buf", "progress_buf",
                          "randomize_buf", "successes", "consecutive_successes", "goal_states", "root_state_tensor",
                          "dof_state", "prev_targets", "cur_targets", "rb_forces", "random_force_prob"]
# sim-owned tensors refreshed by the synthetic "physics" before every step
SIM_INPUTS = {
    "cartpole": ["dof_state"],
    "ant": ["root_states", "dof_state", "vec_sensor_tensor", "actions"],
    "humanoid": ["root_states", "dof_state", "vec_sensor_tensor", "dof_force_tensor", "actions"],
}


SIM_INPUTS["shadow_hand"] = ["dof_state", "root_state_tensor", "rigid_body_states", "vec_sensor_tensor",
                             "dof_force_tensor", "actions"]
OUTPUTS["allegro_hand"], SIM_INPUTS["allegro_hand"] = OUTPUTS["shadow_hand"], SIM_INPUTS["shadow_hand"]
HAND_TASKS = ("shadow_hand", "allegro_hand")
for _v in ("reach", "collect", "combine", "pp"):
    OUTPUTS["franka_" + _v] = ["obs_buf", "rew_buf", "reset_buf", "progress_buf", "_root_state", "_dof_state",
                               "_pos_control", "_effort_control", "_init_cubeA_state"]
    SIM_INPUTS["franka_" + _v] = ["_rigid_body_state", "_root_state", "_dof_state", "_contact_forces", "actions"]


def reset_rand_shapes(task, k, cfg):
    """torch.rand call shapes the reference's reset_idx makes, in order."""
    if task == "cartpole":
        return [(k, 2), (k, 2)]                       # cartpole.py:145-146
    if task in ("ant", "humanoid"):
        return [(k, cfg["num_dof"]), (k, cfg["num_dof"])]   # ant.py:257-258, humanoid.py:258-259
    raise KeyError(task)


def shadow_hand_rand_shapes(k_goal, k_env, cfg):
    """pre_physics_step draw order (shadow_hand.py:670-682): reset_target_pose(goal ids) -> (kg, 4);
    reset_idx -> (k, 2*24+5), its nested reset_target_pose(env ids) -> (k, 4), force prob -> (k,)."""
    shapes = []
    if k_goal > 0:
        shapes.append(("A", (k_goal, 4)))
    if k_env > 0:
        d = cfg["num_hand_dofs"]
        shapes += [("B", (k_env, 2 * d + 5)), ("C", (k_env, 4)), ("D", (k_env,))]
    return shapes


def run_reference_step(task, obj, st, cfg, torch_seed):
    """Copy the sim-owned tensors of ``st`` into the reference object, run one reference
    ``VecTask.step`` and return (outputs dict, reset-rand list, recorded sim calls)."""
    n = obj.progress_buf.shape[0]
    for k in SIM_INPUTS[task]:
        if k == "actions":
            continue
        getattr(obj, k).copy_(_t(st[k]))
    if task.startswith("franka_"):
        return _run_reference_step_ma(task, obj, st, cfg, torch_seed)
    reset_ids = obj.reset_buf.nonzero(as_tuple=False).flatten().clone()
    goal_ids = obj.reset_goal_buf.nonzero(as_tuple=False).flatten().clone() if task in HAND_TASKS else None
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
    if task in HAND_TASKS:
        torch.manual_seed(torch_seed)
        rands = {name: torch.rand(*shape).numpy() for name, shape in
                 shadow_hand_rand_shapes(int(goal_ids.numel()), k, cfg)}
    elif k > 0:
        torch.manual_seed(torch_seed)
        rands = [torch.rand(*s).numpy() for s in reset_rand_shapes(task, k, cfg)]
    out = {name: getattr(obj, name).detach().clone().numpy() for name in OUTPUTS[task]}
    out["obs_clamped"] = obs_dict["obs"].detach().clone().numpy()
    out["time_outs"] = extras["time_outs"].detach().clone().numpy()
    out["reset_env_ids"] = reset_ids.to(torch.int32).numpy()
    if goal_ids is not None:
        out["goal_env_ids"] = goal_ids.to(torch.int32).numpy()
    if "states" in obs_dict:
        out["states_clamped"] = obs_dict["states"].detach().clone().numpy()
    calls = [(name, [a.numpy() if isinstance(a, torch.Tensor) else a for a in args])
             for name, args in obj.gym.calls]
    return out, rands, calls


def _run_reference_step_ma(task, obj, st, cfg, torch_seed):
    a, t, v = cfg["num_agents"], cfg["num_targets"], cfg["variant"]
    for k in SIM_INPUTS[task]:
        if k != "actions":
            getattr(obj, k).copy_(_t(st[k]))
    flags = (obj.reset_buf != 0).view(-1, a).all(dim=1)
    env_ids = flags.nonzero(as_tuple=False).flatten().clone()
    any_flag = bool((obj.reset_buf != 0).any())
    obj.gym.calls.clear()
    torch.manual_seed(torch_seed)
    obs_dict, rew, reset, extras = obj.step(_t(st["actions"]))
    k = int(env_ids.numel())
    rands = []
    if any_flag:      # reset_idx runs (and draws, possibly with k == 0 rows) whenever ANY agent row is flagged
        torch.manual_seed(torch_seed)
        rands = [torch.rand(*s).numpy() for s in franka_rand_shapes(v, k, a, t)]
    out = {name: getattr(obj, name).detach().clone().numpy() for name in OUTPUTS[task]}
    out["obs_clamped"] = obs_dict["obs"].detach().clone().numpy()
    out["time_outs"] = extras["time_outs"].detach().clone().numpy()
    out["reset_env_ids"] = env_ids.to(torch.int32).numpy()
    comps = sorted(k2 for k2 in extras if k2.startswith("r/"))
    if comps:
        out["reward_component_names"] = np.array(comps)
        out["reward_components"] = np.array([float(extras[c]) for c in comps], np.float32)
    calls = [(name, [x.numpy() if isinstance(x, torch.Tensor) else x for x in args]) for name, args in obj.gym.calls]
    return out, rands, calls
