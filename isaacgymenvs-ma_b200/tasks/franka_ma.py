"""Franka multi-agent mirrors: FrankaReachMA / FrankaCollectMA / FrankaCombineMA / FrankaPPMA
(reference: isaacgymenvs/tasks/franka_reach_MA.py, franka_collect_MA.py, franka_combine_MA.py,
franka_PPMA.py).  Buffers carry the fork's flattened agent axis: row = env * num_agents + agent
(franka_reach_MA.py:22-38).  ``post_physics_step`` is one C-ABI call;
``pre_physics_step`` (the OSC torque controller) is the fused ``ige_franka_ma_pre_physics``
kernel once the simulator's mass-matrix / Jacobian tensors are bound with ``attach_dynamics``."""
import numpy as np
import torch

from .. import _abi, _lib
from .base import VecTaskB200, _ptr

COMPONENT_NAMES = ["r/state0", "r/state1", "r/state2", "r/state3", "r/state4", "r/state5", "r/state6", "r/BSR",
                   "r/gBSR", "r/side_rule"]
# which component means each variant logs (franka_collect_MA.py:1160-1168, franka_combine_MA.py:1152-1155,
# franka_PPMA.py:1152-1161); Reach logs none
LOGGED = {"reach": [], "collect": list(range(9)), "combine": [0, 1, 2, 7], "pp": list(range(10))}


class FrankaMA(VecTaskB200):
    ACCUMULATES_STATS = True
    VARIANT = None

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        self.num_agents = int(cfg["num_agents"])          # vec_task.py:102
        self.num_targets = int(cfg["num_targets"])
        self.variant = cfg["variant"]
        if self.variant in ("combine", "pp") and (self.num_agents != 2 or self.num_targets != 2):
            raise ValueError("FrankaCombineMA / FrankaPPMA are written for exactly 2 agents and 2 targets "
                             "(franka_PPMA.py:599-606)")
        if not (1 <= self.num_agents <= _abi.IGE_MA_MAX_AGENTS and 1 <= self.num_targets <= _abi.IGE_MA_MAX_AGENTS):
            raise ValueError("num_agents / num_targets must be in 1..4")
        super().__init__(cfg, sim, device, **kw)
        self.max_episode_length = int(cfg["max_episode_length"])
        t, n, dev = sim.tensors, self.num_envs, self.device
        self._rigid_body_state = t["_rigid_body_state"]
        self._root_state = t["_root_state"]
        self._dof_state = t["_dof_state"]
        self._contact_forces = t["_contact_forces"]
        self._q, self._qd = self._dof_state[..., 0], self._dof_state[..., 1]
        a, tt = self.num_agents, self.num_targets

        def strided(x, start, step, count):
            return x[:, start:start + step * count:step, :]
        self._eef_state = strided(self._rigid_body_state, cfg["eef_body_start"], cfg["eef_body_step"], a)
        self._base_state = strided(self._rigid_body_state, cfg["base_body_start"], cfg["base_body_step"], a)
        self._hands_contact_forces = strided(self._contact_forces, cfg["hand_body_start"], cfg["hand_body_step"], a)
        self._cubeA_state = strided(self._root_state, cfg["cube_actor_start"], cfg["cube_actor_step"], tt)
        self._cubeA_ids = list(range(cfg["cube_actor_start"], cfg["cube_actor_start"] + tt))
        self._franka_ids = list(range(cfg["franka_actor_start"], cfg["franka_actor_start"] + a))
        self._extra_root_ids = []
        if self.variant == "pp":
            self._dest_state = strided(self._root_state, cfg["dest_actor_start"], cfg["dest_actor_step"], tt)
            self._extra_root_ids = list(range(cfg["dest_actor_start"], cfg["dest_actor_start"] + tt))
        if self.variant == "combine":
            self._stack_base_state = self._root_state[:, cfg["stack_base_actor"], :]
            self._extra_root_ids = [cfg["stack_base_actor"]]
        self.num_dofs = self._dof_state.shape[1]
        self.actions = torch.zeros((n * a, self.num_actions), device=dev)
        self.reward_components = torch.zeros(_abi.IGE_MA_NUM_COMPONENTS, device=dev)
        self._reduce_scratch = torch.zeros(_abi.IGE_MA_NUM_COMPONENTS + 2, device=dev)
        self._global_indices = torch.arange(n * cfg["num_actors"], dtype=torch.int32, device=dev).view(n, -1)
        self._u = [None] * 5
        self._p_static = None

    def allocate_buffers(self):
        super().allocate_buffers()      # base already sizes obs/rew/reset/timeout/progress by num_envs * num_agents

    def adopt_state(self, st):
        self._adopt(st, ["_pos_control", "_effort_control", "_init_cubeA_state"])
        for k in ("reset_buf", "progress_buf", "actions"):
            getattr(self, k).copy_(torch.from_numpy(st[k]).to(self.device))

    # ------------------------------------------------------------------------------------------
    def _reset_constants(self):
        """Fold the reset constants exactly like the reference's Python expressions do."""
        c = self.cfg
        f32 = np.float32
        table = np.array(c["table_surface_pos"], dtype=np.float64)   # self._table_surface_pos (numpy, float64)
        size, dsize, noise = c["cubeA_size"], c["dest_size"], c["start_position_noise"]
        out = {}
        if self.variant in ("reach", "collect"):
            # np.float64 + (tensor / 2): fp32 tensor arithmetic (franka_reach_MA.py:683)
            out["cube_z_base"] = float(f32(table[2]) + f32(size) / f32(2.0))
        else:
            # np.float64 + python float first, then cast (franka_PPMA.py:852)
            out["cube_z_base"] = float(f32(table[2] + size / 2))
        out["cube_xy_center"] = [float(f32(table[0])), float(f32(table[1]))]
        out["cube_xy_mult"] = float(f32(2.0 * noise))
        out["aux_z_base"] = float(f32(table[2] + (dsize / 2) * -0.25))          # franka_PPMA.py:880
        if self.variant == "pp":
            cy = float(f32(table[1]) + f32(0.3))
            out["aux_xy_center_x"] = [-0.25, 0.25][:self.num_targets]            # franka_PPMA.py:886-887
            out["aux_xy_center_y"] = cy
            out["aux_xy_mult"] = float(f32(2.0 * noise))
        else:
            out["aux_xy_center_x"] = [float(f32(table[0]) + f32(0.0))]
            out["aux_xy_center_y"] = float(f32(table[1]) + f32(0.0))
            out["aux_xy_mult"] = float(f32(2.0 * 0.15))                          # franka_combine_MA.py:897-900
        out["dof_noise_mult"] = float(f32(c["franka_dof_noise"] * 2.0))
        return out

    def abi_structs(self):
        c = self.cfg
        p = _abi.IgeFrankaMAParams()
        p.variant = _abi.IGE_FRANKA_VARIANT[self.variant]
        p.num_envs, p.num_agents, p.num_targets = self.num_envs, self.num_agents, self.num_targets
        p.num_actions, p.num_obs = self.num_actions, self.num_obs
        p.num_bodies, p.num_actors, p.num_dofs = c["num_bodies"], c["num_actors"], self.num_dofs
        for k in ("eef_body_start", "eef_body_step", "base_body_start", "base_body_step", "hand_body_start",
                  "hand_body_step", "cube_actor_start", "cube_actor_step", "dest_actor_start", "dest_actor_step",
                  "stack_base_actor"):
            setattr(p, k, int(c[k]))
        p.max_episode_length, p.clip_obs = float(self.max_episode_length), self.clip_obs
        p.cubeA_size, p.dest_size = c["cubeA_size"], c["dest_size"]
        p.wall_pos_y, p.wall_height = c["wall_pos_y"], c["wall_height"]
        k = self._reset_constants()
        p.cube_z_base, p.cube_z_range = k["cube_z_base"], c["rand_z_range"]
        _abi.fill_array(p.cube_xy_center, k["cube_xy_center"])
        p.cube_xy_mult, p.aux_z_base = k["cube_xy_mult"], k["aux_z_base"]
        _abi.fill_array(p.aux_xy_center_x, k["aux_xy_center_x"])
        p.aux_xy_center_y, p.aux_xy_mult, p.dof_noise_mult = k["aux_xy_center_y"], k["aux_xy_mult"], k["dof_noise_mult"]
        from ..synthetic import FRANKA_DEFAULT_DOF_POS, FRANKA_DOF_LOWER, FRANKA_DOF_UPPER
        _abi.fill_array(p.franka_default_dof_pos, c.get("franka_default_dof_pos", FRANKA_DEFAULT_DOF_POS))
        _abi.fill_array(p.franka_dof_lower, c.get("franka_dof_lower_limits", FRANKA_DOF_LOWER))
        _abi.fill_array(p.franka_dof_upper, c.get("franka_dof_upper_limits", FRANKA_DOF_UPPER))
        r = p.rand
        if self.rand_mode == "philox":
            r.mode, r.seed, r.step = _abi.IGE_RAND_PHILOX, self.seed, _ptr(self.rng_step)
        else:
            r.mode, r.env_rank = _abi.IGE_RAND_INJECTED, _ptr(self.reset_rank)
            for i in range(5):
                r.u[i] = self._u[i].data_ptr() if self._u[i] is not None else None
        b = _abi.IgeFrankaMABuffers()
        b.rigid_body_state, b.root_state = _ptr(self._rigid_body_state), _ptr(self._root_state)
        b.dof_state, b.contact_forces = _ptr(self._dof_state), _ptr(self._contact_forces)
        b.actions = _ptr(self.actions)
        b.pos_control, b.effort_control = _ptr(self._pos_control), _ptr(self._effort_control)
        b.init_cubeA_state = _ptr(self._init_cubeA_state)
        b.obs_buf, b.rew_buf = _ptr(self.obs_buf), _ptr(self.rew_buf)
        b.reset_buf, b.progress_buf = _ptr(self.reset_buf), _ptr(self.progress_buf)
        b.reward_components = _ptr(self.reward_components)
        b.obs_clamped = _ptr(self.obs_clamped) if self.fuse_epilogue else None
        b.timeout_buf = _ptr(self.timeout_bool) if self.fuse_epilogue else None
        b.stats = _ptr(self.stats_rows)
        b.reduce_scratch = _ptr(self._reduce_scratch)
        b.list = self._list_struct()
        return p, b

    def set_injected_uniforms(self, *u):
        self._u = list(u) + [None] * (5 - len(u))

    def rand_shapes(self, k):
        """torch.rand call shapes of reset_idx in order -> (slot in IgeMAResetRand.u, shape)."""
        t = self.num_targets
        shapes = [(0, (k, t)), (1, (k, t, 2))]
        if self.variant == "pp":
            shapes += [(2, (k, t)), (3, (k, t, 2))]
        if self.variant == "combine":
            shapes += [(2, (k,)), (3, (k, 2))]
        shapes.append((4, (k, 9)))
        return shapes

    def draw_reset_uniforms(self, k):
        u = [None] * 5
        for slot, shape in self.rand_shapes(k):
            u[slot] = torch.rand(shape, device=self.device)
        return u

    # -- pre_physics_step: operational-space torque controller --------------------------------
    EFFORT_LIMITS = [87.0, 87.0, 87.0, 87.0, 12.0, 12.0, 12.0]   # franka_description panda.urdf efforts (:339)
    # controlType "osc" command limits, per axis (franka_reach_MA.py:219, franka_collect_MA.py:228,
    # franka_combine_MA.py / franka_PPMA.py:234)
    CMD_LIMIT = {"reach": [0.1, 0.1, 0.1, 0.5, 0.5, 0.5], "collect": [0.6] * 6, "combine": [0.55] * 6, "pp": [0.55] * 6}

    def attach_dynamics(self, mm, j_eef, action_scale=1.0, cmd_limit=None, kp=150.0, kp_null=10.0,
                        effort_limits=None):
        """Bind the tensors ``_get_mm()`` / ``_get_j_eef()`` return (franka_PPMA.py:1041-1061: PhysX mass matrices
        (N*A, 7, 7) and end-effector Jacobians (N*A, 6, 7), stacked over agents) and the controller gains
        (franka_PPMA.py:226-235).  Afterwards ``pre_physics_step`` is the fused OSC kernel."""
        rows = self.num_envs * self.num_agents
        for name, t, shape in (("mm", mm, (rows, 7, 7)), ("j_eef", j_eef, (rows, 6, 7))):
            if tuple(t.shape) != shape or t.dtype != torch.float32 or t.device != self.device or not t.is_contiguous():
                raise _lib.IgeError(f"{name}: expected a contiguous float32 tensor {shape} on {self.device}")
        if self.num_dofs != 9 * self.num_agents:
            raise _lib.IgeError("OSC pre-physics expects 9 DOFs per Franka (7 arm + 2 fingers)")
        self._mm, self._j_eef = mm, j_eef
        self.action_scale = float(action_scale)
        if cmd_limit is None:
            cmd_limit = self.CMD_LIMIT[self.variant]
        self.cmd_limit = ([float(cmd_limit)] * 6 if isinstance(cmd_limit, (int, float)) else [float(x) for x in cmd_limit])
        if len(self.cmd_limit) != 6:
            raise ValueError("cmd_limit: a scalar or one value per Cartesian axis (6)")
        self.kp = [float(kp)] * 6
        self.kd = (2 * torch.sqrt(torch.tensor(self.kp))).tolist()            # :227
        self.kp_null = [float(kp_null)] * 7
        self.kd_null = (2 * torch.sqrt(torch.tensor(self.kp_null))).tolist()  # :229
        self._franka_effort_limits = list(effort_limits or self.EFFORT_LIMITS)

    def pre_physics_step(self, actions):
        """franka_PPMA.py:931-956 (and the same method of the other three tasks) as one C-ABI call; without
        attached dynamics only ``self.actions = actions.clone()`` (:933) happens."""
        actions = actions.to(self.device)
        if getattr(self, "_mm", None) is None:
            self.actions.copy_(actions)
            return
        if actions.dtype != torch.float32 or not actions.is_contiguous() or tuple(actions.shape) != tuple(self.actions.shape):
            raise _lib.IgeError(f"actions: expected contiguous float32 {tuple(self.actions.shape)}")
        from ..synthetic import FRANKA_DEFAULT_DOF_POS, FRANKA_DOF_LOWER, FRANKA_DOF_UPPER
        c = self.cfg
        a = _abi.IgeFrankaOscArgs()
        a.num_envs, a.num_agents, a.num_actions = self.num_envs, self.num_agents, self.num_actions
        a.has_gripper = int(self.variant != "reach")
        a.num_bodies = self._rigid_body_state.shape[1]
        a.eef_body_start, a.eef_body_step = c["eef_body_start"], c["eef_body_step"]
        a.clip_actions, a.action_scale = self.clip_actions, self.action_scale
        for name, vals in (("cmd_limit", self.cmd_limit), ("kp", self.kp), ("kd", self.kd), ("kp_null", self.kp_null),
                           ("kd_null", self.kd_null), ("effort_limits", self._franka_effort_limits),
                           ("default_dof_pos", list(c.get("franka_default_dof_pos", FRANKA_DEFAULT_DOF_POS))[:7]),
                           ("gripper_lower", list(c.get("franka_dof_lower_limits", FRANKA_DOF_LOWER))[-2:]),
                           ("gripper_upper", list(c.get("franka_dof_upper_limits", FRANKA_DOF_UPPER))[-2:])):
            _abi.fill_array(getattr(a, name), vals)
        a.actions_in, a.actions = _ptr(actions), _ptr(self.actions)
        a.mm, a.j_eef = _ptr(self._mm), _ptr(self._j_eef)
        a.dof_state, a.rigid_body_state = _ptr(self._dof_state), _ptr(self._rigid_body_state)
        a.effort_control, a.pos_control = _ptr(self._effort_control), _ptr(self._pos_control)
        _lib.call("ige_franka_ma_pre_physics", a, stream=self._stream())

    def post_physics_step(self):
        """franka_reach_MA.py:821-828 etc. as one C-ABI call (Combine / PP launch a reset kernel first)."""
        stream = self._stream()
        if self.rand_mode != "philox":
            self.compact_resets(group=self.num_agents)
            any_flag = bool((self.reset_buf != 0).any())   # the reference draws whenever ANY agent row is flagged
            k = int(self.reset_count.item())
            if any_flag and all(u is None for u in self._u):
                self._u = self.draw_reset_uniforms(k)
        p, b = self.abi_structs()
        _lib.call("ige_franka_ma_post_physics", p, b, stream=stream)
        self._u = [None] * 5
        if self.sim.record:
            self.sim_reset_calls()
        for i in LOGGED[self.variant]:
            self.extras[COMPONENT_NAMES[i]] = self.reward_components[i]

    def sim_reset_calls(self):
        """The four gym.set_*_indexed calls of reset_idx (franka_reach_MA.py:676-698): global actor ids of the
        reset envs' frankas and cubes (+ dests / stack base), built from the device-side env id list."""
        k = int(self.reset_count.item())
        env_ids = self.reset_env_ids[:k].long()
        franka_ids = self._global_indices[env_ids][:, self._franka_ids].flatten()
        root_ids = self._global_indices[env_ids][:, self._cubeA_ids + self._extra_root_ids].flatten()
        for name, tensor in (("set_dof_position_target_tensor_indexed", self._pos_control),
                             ("set_dof_actuation_force_tensor_indexed", self._effort_control),
                             ("set_dof_state_tensor_indexed", self._dof_state)):
            self.sim.set_indexed(name, tensor, franka_ids, franka_ids.numel())
        self.sim.set_indexed("set_actor_root_state_tensor_indexed", self._root_state, root_ids, root_ids.numel())


class FrankaReachMA(FrankaMA):
    VARIANT = "reach"


class FrankaCollectMA(FrankaMA):
    VARIANT = "collect"


class FrankaCombineMA(FrankaMA):
    VARIANT = "combine"


class FrankaPPMA(FrankaMA):
    VARIANT = "pp"
