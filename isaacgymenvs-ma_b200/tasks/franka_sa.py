"""Single-agent Franka mirrors: FrankaReach, FrankaCubeStack2, FrankaCubeStack and
FrankaCabinet (reference: isaacgymenvs/tasks/franka_reach.py, franka_cube_stack2.py, franka_cube_stack.py,
franka_cabinet.py).

``post_physics_step`` is ONE C-ABI call (``ige_franka_sa_post_physics`` / ``ige_franka_cabinet_post_physics``):
progress++, ``reset_idx`` of the flagged envs (cube re-sampling, arm joints), ``_update_states`` + FSM, observations,
``compute_franka_reward`` and the ``step()`` epilogue; the kernel also emits the ascending env id list the reference
hands to ``gym.set_*_tensor_indexed``.  Attribute names and tensor layouts follow the reference.  No CPU fallback."""
import math

import numpy as np
import torch

from .. import _abi, _lib, synthetic
from .base import VecTaskB200, _ptr

F32 = np.float32
VARIANT_ID = {"reach": _abi.IGE_FRANKA_SA_REACH, "stack2": _abi.IGE_FRANKA_SA_STACK2, "stack": _abi.IGE_FRANKA_SA_STACK}
# slot of each torch.rand draw of reset_idx in rand.u[] (reference call order), -1 = not drawn by the variant:
#   [0] cubeB z  [1] cubeB xy  [2] cubeB yaw  [3] cubeA z  [4] cubeA xy  [5] cubeA yaw  [6] franka joints
DRAW_SLOT = {"reach": (-1, -1, -1, 0, 1, 2, 3),              # franka_reach.py:484, :596-629, :492
             "stack2": (0, 1, 2, 3, 4, 5, 6),                # franka_cube_stack2.py:506-507, :603-644, :514
             "stack": (-1, 0, 1, -1, 2, 3, 4)}               # franka_cube_stack.py:462-463 (no z draw), :567-597, :470
DRAW_COLS = {"reach": (1, 2, 1, 9), "stack2": (1, 2, 1, 1, 2, 1, 9), "stack": (2, 1, 2, 1, 9)}
SA_TENSORS = dict(rigid_body_state=None, root_state=None, dof_state=(9, 2), actions=(7,), pos_control=(9,),
                  effort_control=(9,), init_cubeA_state=(13,), init_cubeB_state=(13,))
COMPONENT_NAMES = {"reach": {"r/state0": 0, "r/state1": 1, "r/state2": 2, "r/state": 7, "mean_cube_height": 8},
                   "stack2": {"r/state0": 0, "r/state1": 1, "r/state2": 2, "r/stat3": 3, "r/stat4": 4, "r/stat5": 5,
                              "r/state6": 6, "r/state": 7, "mean_cube_height": 8, "target_dist": 9},
                   "stack": {}}


def fill_sa_scalars(a, cfg, num_envs):
    """cfg (synthetic.franka_sa_cfg) -> scalar fields of IgeFrankaSAArgs.  Python-float products are formed in double
    precision and rounded once; tensor-with-tensor products of the reference are formed in fp32, as there."""
    v = cfg["variant"]
    a.variant, a.num_envs = VARIANT_ID[v], num_envs
    a.num_bodies, a.num_actors = cfg["num_bodies"], cfg["num_actors"]
    a.eef_body, a.lf_body, a.rf_body = cfg["eef_body"], cfg["lf_body"], cfg["rf_body"]
    a.cubeA_actor, a.cubeB_actor = cfg["cubeA_actor"], cfg["cubeB_actor"]
    _abi.fill_array(a.draw_slot, DRAW_SLOT[v])
    a.max_episode_length, a.clip_obs = float(cfg["max_episode_length"]), float(cfg.get("clip_obs", math.inf))
    sa, sb = F32(cfg["cubeA_size"]), F32(cfg["cubeB_size"])
    table = cfg["table_surface_pos"]
    a.cubeA_size, a.cubeB_size, a.table_height = float(sa), float(sb), float(table[2])
    _abi.fill_array(a.cube_xy_center, table[:2])
    a.cube_xy_mult = 2.0 * cfg["start_position_noise"]
    # `table_z + cube_heights / 2 (+ rand * range)` (franka_reach.py:598): fp32 tensor arithmetic
    a.cubeA_z_base = float(F32(table[2]) + sa / F32(2))
    a.cubeB_z_base = float(F32(table[2]) + (sa if v == "stack" else sb) / F32(2))   # franka_cube_stack.py:537 uses cubeA's
    a.cube_z_range = float(cfg["rand_z_range"])
    a.cube_yaw_mult = 2.0 * cfg["start_rotation_noise"]
    a.dof_noise_mult = cfg["franka_dof_noise"] * 2.0
    a.min_dist = float((sa + sb) * F32(math.sqrt(2)) / F32(2) * F32(2))            # franka_cube_stack.py:550-553
    for k in ("r_dist_scale", "r_lift_scale", "r_align_scale", "r_stack_scale"):
        setattr(a, k, float(cfg[k]))
    _abi.fill_array(a.default_dof_pos, synthetic.FRANKA_DEFAULT_DOF_POS)
    _abi.fill_array(a.dof_lower, synthetic.FRANKA_DOF_LOWER)
    _abi.fill_array(a.dof_upper, synthetic.FRANKA_DOF_UPPER)
    return a


def fill_cabinet_scalars(a, cfg, num_envs):
    a.num_envs = num_envs
    for k in ("num_bodies", "num_actors", "num_dofs", "num_props", "hand_body", "drawer_body", "lfinger_body",
              "rfinger_body"):
        setattr(a, k, int(cfg[k]))
    a.max_episode_length, a.clip_obs = float(cfg["max_episode_length"]), float(cfg.get("clip_obs", math.inf))
    a.target_step = float(cfg["dt"])
    for k in ("dof_vel_scale", "action_scale", "dist_reward_scale", "rot_reward_scale", "around_handle_reward_scale",
              "open_reward_scale", "finger_dist_reward_scale", "action_penalty_scale", "distX_offset"):
        setattr(a, k, float(cfg[k]))
    for k in ("franka_local_grasp_pos", "franka_local_grasp_rot", "drawer_local_grasp_pos", "drawer_local_grasp_rot",
              "gripper_forward_axis", "drawer_inward_axis", "gripper_up_axis", "drawer_up_axis", "default_dof_pos",
              "dof_speed_scales"):
        _abi.fill_array(getattr(a, k), cfg[k])
    _abi.fill_array(a.dof_lower, synthetic.FRANKA_DOF_LOWER)
    _abi.fill_array(a.dof_upper, synthetic.FRANKA_DOF_UPPER)
    return a


class _PostResetTask(VecTaskB200):
    """Shared plumbing: injected draws by reset rank (parity) or in-kernel Philox (production), one launch."""
    ENTRY = None
    DRAWS = ()

    def _common_ptrs(self, a):
        a.obs_buf, a.rew_buf = _ptr(self.obs_buf), _ptr(self.rew_buf)
        a.reset_buf, a.progress_buf = _ptr(self.reset_buf), _ptr(self.progress_buf)
        a.obs_clamped = _ptr(self.obs_clamped) if self.fuse_epilogue else None
        a.timeout_buf = _ptr(self.timeout_bool) if self.fuse_epilogue else None
        a.list = self._list_struct()

    def _rand_ptrs(self, a):
        if self.rand_mode == "philox":
            a.rand.mode, a.rand.seed, a.rand.step = _abi.IGE_RAND_PHILOX, self.seed, _ptr(self.rng_step)
        else:
            a.rand.mode, a.rand.rank = _abi.IGE_RAND_INJECTED, _ptr(self.reset_rank)
            for i, u in enumerate(self._u):
                a.rand.u[i] = u.data_ptr() if u is not None and u.numel() else None

    def draw_reset_uniforms(self, k):
        """INJECTED mode: the torch.rand calls of reset_idx, reference order and shapes."""
        return [torch.rand((k, c), device=self.device) if k > 0 else None for c in self.DRAWS]

    def _on_remap(self):
        self._args = None

    def post_physics_step(self):
        if self.rand_mode != "philox":
            self.compact_resets()
            k = int(self.reset_count.item())            # the reference's own host sync: len(env_ids)
            if all(u is None for u in self._u):
                self._u = self.draw_reset_uniforms(k)
        _lib.call(self.ENTRY, self.abi_struct(), stream=self._stream())
        self._u = [None] * len(self.DRAWS)
        self._apply_sim_calls()


class FrankaSA(_PostResetTask):
    """FrankaReach / FrankaCubeStack2 / FrankaCubeStack; subclasses name the variant."""
    VARIANT = None
    ENTRY = "ige_franka_sa_post_physics"

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        super().__init__(cfg, sim, device, **kw)
        n, dev = self.num_envs, self.device
        self.DRAWS = DRAW_COLS[self.VARIANT]
        self.max_episode_length = int(cfg["max_episode_length"])
        self._rigid_body_state = sim.tensors["rigid_body_state"].view(n, cfg["num_bodies"], 13)   # franka_reach.py:366-376
        self._root_state = sim.tensors["root_state"].view(n, cfg["num_actors"], 13)
        self._dof_state = sim.tensors["dof_state"].view(n, 9, 2)
        self._q, self._qd = self._dof_state[..., 0], self._dof_state[..., 1]
        self._eef_state = self._rigid_body_state[:, cfg["eef_body"], :]
        self._cubeA_state = self._root_state[:, cfg["cubeA_actor"], :]
        self._cubeB_state = self._root_state[:, cfg["cubeB_actor"], :]
        self.actions = torch.zeros((n, 7), device=dev)
        self._pos_control = torch.zeros((n, 9), device=dev)                                      # :399-404
        self._effort_control = torch.zeros_like(self._pos_control)
        self._init_cubeA_state = torch.zeros((n, 13), device=dev)
        self._init_cubeB_state = torch.zeros((n, 13), device=dev)
        self._global_indices = torch.arange(n * cfg["num_actors"], dtype=torch.int32, device=dev).view(n, -1)
        self.reward_components = torch.zeros(_abi.IGE_FRANKA_SA_COMPONENTS, device=dev)
        self._reduce_scratch = torch.zeros(_abi.IGE_FRANKA_SA_COMPONENTS + 2, device=dev)
        self._u = [None] * len(self.DRAWS)
        self._args = None

    def adopt_state(self, st):
        for k in ("reset_buf", "progress_buf"):
            getattr(self, k).copy_(torch.from_numpy(st[k]).to(self.device))
        for k, attr in (("actions", "actions"), ("pos_control", "_pos_control"), ("effort_control", "_effort_control"),
                        ("init_cubeA_state", "_init_cubeA_state"), ("init_cubeB_state", "_init_cubeB_state")):
            if k in st:
                getattr(self, attr).copy_(torch.from_numpy(st[k]).to(self.device))
        self.actions.clamp_(-self.clip_actions, self.clip_actions)                               # vec_task.py:373

    def abi_struct(self):
        a = self._args
        if a is None:
            a = self._args = fill_sa_scalars(_abi.IgeFrankaSAArgs(), self.cfg, self.num_envs)
            a.rigid_body_state, a.root_state, a.dof_state = (_ptr(self._rigid_body_state), _ptr(self._root_state),
                                                             _ptr(self._dof_state))
            a.actions, a.pos_control, a.effort_control = _ptr(self.actions), _ptr(self._pos_control), _ptr(self._effort_control)
            a.init_cubeA_state, a.init_cubeB_state = _ptr(self._init_cubeA_state), _ptr(self._init_cubeB_state)
            a.reward_components, a.reduce_scratch = _ptr(self.reward_components), _ptr(self._reduce_scratch)
            self._common_ptrs(a)
        self._rand_ptrs(a)
        return a

    def pre_physics_step(self, actions):
        """The OSC torque controller (franka_reach.py:635-693) needs PhysX Jacobians / mass matrices; its MA twin is
        ``ige_franka_ma_pre_physics``.  Here: ``self.actions = actions.clone()`` (:664)."""
        self.actions.copy_(actions.to(self.device).view_as(self.actions))

    def _apply_sim_calls(self):
        ids, cnt = self.reset_env_ids, self.reset_count                                           # :507-535
        self.sim.set_indexed("set_dof_position_target_tensor_indexed", self._pos_control, ids, cnt)
        self.sim.set_indexed("set_dof_actuation_force_tensor_indexed", self._effort_control, ids, cnt)
        self.sim.set_indexed("set_dof_state_tensor_indexed", self._dof_state, ids, cnt)
        self.sim.set_indexed("set_actor_root_state_tensor_indexed", self._root_state, ids, cnt)

    @property
    def extras_components(self):
        """``self.extras.update(self.reward_components)`` + mean_cube_height / target_dist (franka_reach.py:735-738,
        franka_cube_stack2.py:745-757) as a dict of 0-d tensors (no host sync)."""
        return {k: self.reward_components[i] for k, i in COMPONENT_NAMES[self.VARIANT].items()}


class FrankaReach(FrankaSA):
    VARIANT = "reach"


class FrankaCubeStack2(FrankaSA):
    VARIANT = "stack2"


class FrankaCubeStack(FrankaSA):
    VARIANT = "stack"


class FrankaCabinet(_PostResetTask):
    ENTRY = "ige_franka_cabinet_post_physics"
    DRAWS = (9,)

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        super().__init__(cfg, sim, device, **kw)
        n, dev, D = self.num_envs, self.device, cfg["num_dofs"]
        self.max_episode_length = int(cfg["max_episode_length"])
        self.dt, self.action_scale = float(cfg["dt"]), float(cfg["action_scale"])
        self.num_franka_dofs, self.num_props = 9, int(cfg["num_props"])
        self.rigid_body_states = sim.tensors["rigid_body_states"].view(n, cfg["num_bodies"], 13)  # franka_cabinet.py:93-113
        self.root_state_tensor = sim.tensors["root_state_tensor"].view(n, cfg["num_actors"], 13)
        self.dof_state = sim.tensors["dof_state"].view(n, D, 2)
        self.franka_dof_state = self.dof_state[:, :9]
        self.franka_dof_pos, self.franka_dof_vel = self.franka_dof_state[..., 0], self.franka_dof_state[..., 1]
        self.cabinet_dof_state = self.dof_state[:, 9:]
        self.cabinet_dof_pos, self.cabinet_dof_vel = self.cabinet_dof_state[..., 0], self.cabinet_dof_state[..., 1]
        self.prop_states = self.root_state_tensor[:, 2:]
        self.default_prop_states = torch.zeros((n, self.num_props, 13), device=dev)
        self.franka_dof_targets = torch.zeros((n, D), device=dev)
        self.actions = torch.zeros((n, 9), device=dev)
        self.franka_grasp_pos, self.franka_grasp_rot = torch.zeros((n, 3), device=dev), torch.zeros((n, 4), device=dev)
        self.drawer_grasp_pos, self.drawer_grasp_rot = torch.zeros((n, 3), device=dev), torch.zeros((n, 4), device=dev)
        self.global_indices = torch.arange(n * cfg["num_actors"], dtype=torch.int32, device=dev).view(n, -1)
        self._u = [None]
        self._args = None

    def adopt_state(self, st):
        for k in ("reset_buf", "progress_buf", "default_prop_states", "franka_dof_targets", "actions"):
            if k in st:
                getattr(self, k).copy_(torch.from_numpy(st[k]).to(self.device))
        self.actions.clamp_(-self.clip_actions, self.clip_actions)

    def abi_struct(self):
        a = self._args
        if a is None:
            a = self._args = fill_cabinet_scalars(_abi.IgeFrankaCabinetArgs(), self.cfg, self.num_envs)
            for k in ("rigid_body_states", "root_state_tensor", "default_prop_states", "dof_state", "actions",
                      "franka_dof_targets", "franka_grasp_pos", "franka_grasp_rot", "drawer_grasp_pos", "drawer_grasp_rot"):
                setattr(a, k, _ptr(getattr(self, k)))
            self._common_ptrs(a)
        self._rand_ptrs(a)
        return a

    def pre_physics_step(self, actions):                                                         # :417-425
        self.actions.copy_(actions.to(self.device).view_as(self.actions))
        _lib.call("ige_franka_cabinet_pre_physics", self.abi_struct(), stream=self._stream())
        self.sim.set_indexed("set_dof_position_target_tensor", self.franka_dof_targets, None, None)

    def _apply_sim_calls(self):
        ids, cnt = self.reset_env_ids, self.reset_count                                           # :401-412
        if self.num_props > 0:
            self.sim.set_indexed("set_actor_root_state_tensor_indexed", self.root_state_tensor, ids, cnt)
        self.sim.set_indexed("set_dof_position_target_tensor_indexed", self.franka_dof_targets, ids, cnt)
        self.sim.set_indexed("set_dof_state_tensor_indexed", self.dof_state, ids, cnt)
