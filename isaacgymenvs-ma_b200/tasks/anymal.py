"""Anymal mirror (reference: isaacgymenvs/tasks/anymal.py) -- SURVEY.md 8f row 2.

Same attribute names and buffer contracts as the reference class; ``post_physics_step`` (:231-240) is ONE C-ABI call
(``ige_anymal_post_physics``) that also emits the reset index list the two ``set_*_indexed`` calls of ``reset_idx``
consume (:290-296).  No CPU fallback."""
import torch

from .. import _abi, _lib
from .base import VecTaskB200, _ptr


class Anymal(VecTaskB200):
    FUSED_ACTION_CLAMP = False

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        cfg.setdefault("num_obs", 48)       # anymal.py:88
        cfg.setdefault("num_actions", 12)   # anymal.py:89
        super().__init__(cfg, sim, device, **kw)
        n, dev = self.num_envs, self.device
        for k in ("lin_vel_scale", "ang_vel_scale", "dof_pos_scale", "dof_vel_scale", "action_scale"):
            setattr(self, k, float(cfg[k]))                                       # anymal.py:49-53
        self.dt = float(cfg.get("dt", 0.02))
        self.rew_scales = {"lin_vel_xy": cfg["rew_lin_vel_xy"] * self.dt, "ang_vel_z": cfg["rew_ang_vel_z"] * self.dt,
                           "torque": cfg["rew_torque"] * self.dt}                 # anymal.py:56-59, :99-100
        self.command_x_range = list(cfg["command_x_range"])                       # anymal.py:66-68
        self.command_y_range = list(cfg["command_y_range"])
        self.command_yaw_range = list(cfg["command_yaw_range"])
        self.max_episode_length = int(cfg["max_episode_length"])                  # anymal.py:95
        self.num_dof, self.num_bodies = 12, int(cfg["num_bodies"])
        self.base_index = int(cfg["base_index"])
        self._knees = [int(k) for k in cfg["knee_indices"]]      # host copy: abi_structs() must not sync (graph capture)
        self.knee_indices = torch.tensor(self._knees, dtype=torch.long, device=dev)
        t = sim.tensors
        self.root_states, self.dof_state = t["root_states"], t["dof_state"]       # anymal.py:121-126
        self.dof_pos = self.dof_state.view(n, self.num_dof, 2)[..., 0]
        self.dof_vel = self.dof_state.view(n, self.num_dof, 2)[..., 1]
        self.contact_forces, self.torques = t["contact_forces"].view(n, -1, 3), t["torques"].view(n, self.num_dof)
        self.commands = torch.zeros((n, 3), device=dev)                           # anymal.py:128-131
        self.commands_x, self.commands_y, self.commands_yaw = (self.commands.view(n, 3)[..., i] for i in range(3))
        self._default_row = [float(x) for x in cfg["default_dof_pos"]]
        self.default_dof_pos = torch.tensor(self._default_row, device=dev).repeat(n, 1)   # anymal.py:132-138
        self.initial_root_states = self.root_states.clone()
        self.gravity_vec = torch.tensor([0.0, 0.0, -1.0], device=dev).repeat(n, 1)
        self.actions = torch.zeros((n, self.num_actions), device=dev)
        words = _lib.lib().ige_anymal_scratch_words(max(n, 1))
        self._anymal_scratch = torch.zeros((words,), device=dev, dtype=torch.int32)
        self._u = None

    def adopt_state(self, st):
        for k in ("reset_buf", "progress_buf", "commands"):
            getattr(self, k).copy_(torch.from_numpy(st[k]).to(self.device))
        self.actions.copy_(torch.clamp(torch.from_numpy(st["actions"]), -self.clip_actions, self.clip_actions).to(self.device))

    def draw_reset_uniforms(self, k):
        """The five torch.rand calls of reset_idx, in the reference's order and shapes (anymal.py:282-302)."""
        return [torch.rand(s, device=self.device) for s in ((k, 12), (k, 12), (k, 1), (k, 1), (k, 1))]

    def abi_structs(self):
        p = _abi.IgeAnymalParams()
        p.num_envs, p.num_bodies, p.base_index = self.num_envs, self.num_bodies, self.base_index
        knees = self._knees
        p.num_knees = len(knees)
        _abi.fill_array(p.knee_indices, knees)
        p.lin_vel_scale, p.ang_vel_scale = self.lin_vel_scale, self.ang_vel_scale
        p.dof_pos_scale, p.dof_vel_scale = self.dof_pos_scale, self.dof_vel_scale
        p.rew_lin_vel_xy, p.rew_ang_vel_z = self.rew_scales["lin_vel_xy"], self.rew_scales["ang_vel_z"]
        p.rew_torque = self.rew_scales["torque"]
        p.max_episode_length, p.clip_obs = float(self.max_episode_length), self.clip_obs
        _abi.fill_array(p.command_x_range, self.command_x_range)
        _abi.fill_array(p.command_y_range, self.command_y_range)
        _abi.fill_array(p.command_yaw_range, self.command_yaw_range)
        _abi.fill_array(p.default_dof_pos, self._default_row)
        if self.rand_mode == "philox":
            p.rand_mode, p.seed, p.step = _abi.IGE_RAND_PHILOX, self.seed, _ptr(self.rng_step)
        else:
            p.rand_mode = _abi.IGE_RAND_INJECTED
            for i, u in enumerate(self._u or []):
                p.u[i] = u.data_ptr() if u.numel() else None
        b = _abi.IgeAnymalBuffers()
        b.root_states, b.dof_state = _ptr(self.root_states), _ptr(self.dof_state)
        b.contact_forces, b.torques = _ptr(self.contact_forces), _ptr(self.torques)
        b.actions, b.commands = _ptr(self.actions), _ptr(self.commands)
        b.obs_buf, b.rew_buf = _ptr(self.obs_buf), _ptr(self.rew_buf)
        b.reset_buf, b.progress_buf = _ptr(self.reset_buf), _ptr(self.progress_buf)
        b.obs_clamped = _ptr(self.obs_clamped) if self.fuse_epilogue else None
        b.timeout_buf = _ptr(self.timeout_bool) if self.fuse_epilogue else None
        b.reset_ids, b.reset_count = _ptr(self.reset_env_ids), _ptr(self.reset_count)
        b.scratch = _ptr(self._anymal_scratch)
        return p, b

    def pre_physics_step(self, actions):
        # anymal.py:226-229 (position targets; element-wise, not on the post-physics path)
        self.actions.copy_(actions.to(self.device))
        targets = self.action_scale * self.actions + self.default_dof_pos
        self.sim.set_indexed("set_dof_position_target_tensor", targets, None, None)

    def post_physics_step(self):
        """anymal.py:231-240 as one fused kernel; the reset list feeds the two indexed sim calls of :290-296."""
        stream = self._stream()
        if self.rand_mode != "philox" and self._u is None:
            k = int((self.reset_buf != 0).sum())      # the reference's own host sync: len(env_ids), anymal.py:235
            self._u = self.draw_reset_uniforms(k)
        p, b = self.abi_structs()
        _lib.call("ige_anymal_post_physics", p, b, stream=stream)
        self.sim.set_indexed("set_actor_root_state_tensor_indexed", self.initial_root_states, self.reset_env_ids,
                             self.reset_count)
        self.sim.set_indexed("set_dof_state_tensor_indexed", self.dof_state, self.reset_env_ids, self.reset_count)
        self._u = None
