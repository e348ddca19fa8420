"""Ant / Humanoid mirrors (reference: isaacgymenvs/tasks/ant.py, isaacgymenvs/tasks/humanoid.py)."""
import torch

from .. import _abi, _lib
from .base import VecTaskB200, _ptr


class _Locomotion(VecTaskB200):
    ACCUMULATES_STATS = True
    ENTRY = None
    HAS_DOF_FORCE = False
    SENSORS = 0
    HOST_SIM_ATTRS = ("root_states", "dof_state", "vec_sensor_tensor", "dof_force_tensor", "actions")
    HOST_INOUT_ATTRS = ("dof_state",)
    FUSED_ACTION_CLAMP = True

    def _on_remap(self):
        n, d = self.num_envs, self.num_dof
        self.dof_pos = self.dof_state.view(n, d, 2)[..., 0]
        self.dof_vel = self.dof_state.view(n, d, 2)[..., 1]

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        cfg.setdefault("num_actions", cfg["num_dof"])
        super().__init__(cfg, sim, device, **kw)
        for k in ("dof_vel_scale", "contact_force_scale", "power_scale", "heading_weight", "up_weight",
                  "actions_cost_scale", "energy_cost_scale", "joints_at_limit_cost_scale", "death_cost",
                  "termination_height", "dt"):
            setattr(self, k, float(cfg[k]))
        self.angular_velocity_scale = float(cfg.get("angular_velocity_scale", 0.1))  # humanoid.py:49
        self.max_episode_length = int(cfg["max_episode_length"])
        self.num_dof = int(cfg["num_dof"])
        self.up_axis_idx = 2
        self.potentials_recip_mul = int(cfg.get("potentials_recip_mul", 0))
        t = sim.tensors
        n, d = self.num_envs, self.num_dof
        # gymtorch views (ant.py:78-95)
        self.root_states = t["root_states"]
        self.dof_state = t["dof_state"]
        self.vec_sensor_tensor = t["vec_sensor_tensor"]
        self.dof_force_tensor = t.get("dof_force_tensor")
        self.dof_pos = self.dof_state.view(n, d, 2)[..., 0]
        self.dof_vel = self.dof_state.view(n, d, 2)[..., 1]
        self.actions = torch.zeros((n, d), device=self.device)
        self.forces = None

    def adopt_state(self, st):
        """Task-owned tensors that the reference builds in __init__ (ant.py:88-114)."""
        self._adopt(st, ["initial_root_states", "initial_dof_pos", "dof_limits_lower", "dof_limits_upper",
                         "targets", "inv_start_rot", "basis_vec0", "basis_vec1", "potentials",
                         "prev_potentials", "up_vec", "heading_vec"])
        for k in ("reset_buf", "progress_buf", "randomize_buf", "actions"):
            if k in st:
                getattr(self, k).copy_(torch.from_numpy(st[k]).to(self.device))
        if "motor_efforts" in st:
            self._adopt(st, ["motor_efforts"])
            self.max_motor_effort = float(self.motor_efforts.max())
        self._consts = None

    def _constants(self):
        if self._consts is None:
            c = {k: self._row_constant(getattr(self, k), k)
                 for k in ("targets", "inv_start_rot", "basis_vec0", "basis_vec1")}
            c["lo"] = self.dof_limits_lower.tolist()
            c["hi"] = self.dof_limits_upper.tolist()
            if hasattr(self, "motor_efforts"):
                # humanoid.py:355.  Divided on the host: torch-CUDA turns `tensor / python_float` into a
                # multiply by the reciprocal (1/3 -> 0.33333331 instead of 0.33333334), which would move
                # the reward by an ulp against the CPU-pinned reference.
                c["ratio"] = (self.motor_efforts.detach().cpu() / self.max_motor_effort).tolist()
            self._consts = c
        return self._consts

    def abi_structs(self):
        c = self._constants()
        p = _abi.IgeLocomotionParams()
        p.num_envs, p.num_dof, p.num_sensor_floats = self.num_envs, self.num_dof, self.SENSORS * 6
        p.potentials_recip_mul = self.potentials_recip_mul
        p.dt = self.dt
        p.dof_vel_scale = self.dof_vel_scale
        p.contact_force_scale = self.contact_force_scale
        p.angular_velocity_scale = self.angular_velocity_scale
        p.up_weight, p.heading_weight = self.up_weight, self.heading_weight
        p.actions_cost_scale = self.actions_cost_scale
        p.energy_cost_scale = self.energy_cost_scale
        p.joints_at_limit_cost_scale = self.joints_at_limit_cost_scale
        p.termination_height, p.death_cost = self.termination_height, self.death_cost
        p.max_episode_length = float(self.max_episode_length)
        p.clip_obs = self.clip_obs
        _abi.fill_array(p.targets, c["targets"])
        _abi.fill_array(p.inv_start_rot, c["inv_start_rot"])
        _abi.fill_array(p.basis_vec0, c["basis_vec0"])
        _abi.fill_array(p.basis_vec1, c["basis_vec1"])
        _abi.fill_array(p.dof_limits_lower, c["lo"])
        _abi.fill_array(p.dof_limits_upper, c["hi"])
        if "ratio" in c:
            _abi.fill_array(p.motor_effort_ratio, c["ratio"])
        p.rand = self._rand_struct()
        b = _abi.IgeLocomotionBuffers()
        for name in ("root_states", "dof_state", "vec_sensor_tensor", "actions", "initial_root_states",
                     "initial_dof_pos", "potentials", "prev_potentials", "up_vec", "heading_vec", "obs_buf",
                     "rew_buf", "reset_buf", "progress_buf", "randomize_buf"):
            setattr(b, name, _ptr(getattr(self, name)))
        b.dof_force_tensor = _ptr(self.dof_force_tensor) if self.HAS_DOF_FORCE else None
        b.obs_clamped = _ptr(self.obs_clamped) if self.fuse_epilogue else None
        b.timeout_buf = _ptr(self.timeout_bool) if self.fuse_epilogue else None
        b.stats = _ptr(self.stats_rows)
        b.list = self._list_struct()
        return p, b

    def draw_reset_uniforms(self, k):
        # ant.py:257-258: torch_rand_float twice, positions then velocities, shape (k, num_dof)
        u0 = torch.rand((k, self.num_dof), device=self.device)
        u1 = torch.rand((k, self.num_dof), device=self.device)
        return u0, u1

    def pre_physics_step(self, actions):
        """ant.py:281-285 / humanoid.py:281-285 (+ the clamp of VecTask.step, vec_task.py:373) as one launch:
        self.actions = clamp(actions), forces = actions * gears * power_scale."""
        if self.forces is None:
            self.forces = torch.zeros_like(self.actions)
        src = actions.to(self.device, dtype=torch.float32).contiguous()
        if src.data_ptr() == self.actions.data_ptr():
            src = src.clone()
        a = _abi.IgeActionForceArgs()
        a.num_envs, a.num_dof = self.num_envs, self.num_dof
        a.clip_actions, a.power_scale = self.clip_actions, self.power_scale
        gears = getattr(self, "motor_efforts", None)
        _abi.fill_array(a.gears, gears.tolist() if gears is not None else [15.0] * self.num_dof)  # ant.py joint_gears
        a.actions_in, a.actions, a.forces = _ptr(src), _ptr(self.actions), _ptr(self.forces)
        _lib.call("ige_locomotion_pre_physics", a, stream=self._stream())
        self._actions_in = src          # keep the source alive until the stream has consumed it
        self.sim.set_indexed("set_dof_actuation_force_tensor", self.forces, None, None)

    def post_physics_step(self):
        """ant.py:287-297 / humanoid.py:287-296 as: device-side nonzero() + ONE fused kernel."""
        stream = self._stream()
        if self.rand_mode != "philox":
            # parity mode: the uniforms are drawn on the host side in the reference's order, which needs
            # len(env_ids) first -- the same host sync the reference has
            self.compact_resets()
            k = int(self.reset_count.item())          # the reference's own host sync (ant.py:292)
            if k > 0 and self._u0 is None:
                self._u0, self._u1 = self.draw_reset_uniforms(k)
        p, b = self.abi_structs()
        _lib.call(self.ENTRY, p, b, stream=stream)
        # ant.py:265-271: the two indexed sim writes (ids int32 ascending + device count)
        self.sim.set_indexed("set_actor_root_state_tensor_indexed", self.initial_root_states,
                             self.reset_env_ids, self.reset_count)
        self.sim.set_indexed("set_dof_state_tensor_indexed", self.dof_state, self.reset_env_ids, self.reset_count)
        self.extras["true_objective"] = self.root_states[:, 7]   # ant.py:245-250
        self._u0 = self._u1 = None


class Ant(_Locomotion):
    ENTRY = "ige_ant_post_physics"
    SENSORS = 4

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        cfg.setdefault("num_obs", 60)   # ant.py:67
        cfg.setdefault("num_dof", 8)
        super().__init__(cfg, sim, device, **kw)


class Humanoid(_Locomotion):
    ENTRY = "ige_humanoid_post_physics"
    SENSORS = 2
    HAS_DOF_FORCE = True

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        cfg.setdefault("num_obs", 108)  # humanoid.py:66
        cfg.setdefault("num_dof", 21)
        super().__init__(cfg, sim, device, **kw)

    def post_physics_step(self):
        super().post_physics_step()
        self.extras.pop("true_objective", None)   # Humanoid has no compute_true_objective
