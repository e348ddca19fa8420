"""Cartpole mirror (reference: isaacgymenvs/tasks/cartpole.py)."""
import torch

from .. import _abi, _lib
from .base import VecTaskB200, _ptr


class Cartpole(VecTaskB200):
    ACCUMULATES_STATS = True
    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        cfg.setdefault("num_obs", 4)          # cartpole.py:46
        cfg.setdefault("num_actions", 1)      # cartpole.py:47
        super().__init__(cfg, sim, device, **kw)
        self.reset_dist = float(cfg["reset_dist"])              # cartpole.py:41
        self.max_push_effort = float(cfg.get("max_effort", 400.0))
        self.max_episode_length = int(cfg.get("max_episode_length", 500))  # cartpole.py:44
        self.num_dof = 2
        self.dof_state = sim.tensors["dof_state"]               # cartpole.py:51-54
        self.dof_pos = self.dof_state.view(self.num_envs, self.num_dof, 2)[..., 0]
        self.dof_vel = self.dof_state.view(self.num_envs, self.num_dof, 2)[..., 1]

    def abi_structs(self):
        p = _abi.IgeCartpoleParams()
        p.num_envs = self.num_envs
        p.reset_dist = self.reset_dist
        p.max_episode_length = float(self.max_episode_length)
        p.clip_obs = self.clip_obs
        p.rand = self._rand_struct()
        b = _abi.IgeCartpoleBuffers()
        b.dof_state = _ptr(self.dof_state)
        b.obs_buf = _ptr(self.obs_buf)
        b.rew_buf = _ptr(self.rew_buf)
        b.reset_buf = _ptr(self.reset_buf)
        b.progress_buf = _ptr(self.progress_buf)
        b.obs_clamped = _ptr(self.obs_clamped) if self.fuse_epilogue else None
        b.timeout_buf = _ptr(self.timeout_bool) if self.fuse_epilogue else None
        b.stats = _ptr(self.stats_rows)
        b.list = self._list_struct()
        return p, b

    def draw_reset_uniforms(self, k):
        # cartpole.py:145-146: two torch.rand((k, num_dof)) draws, positions then velocities
        u0 = torch.rand((k, self.num_dof), device=self.device)
        u1 = torch.rand((k, self.num_dof), device=self.device)
        return u0, u1

    def pre_physics_step(self, actions):
        # cartpole.py:159-163 (out of the hot path; plain torch)
        actions_tensor = torch.zeros(self.num_envs * self.num_dof, device=self.device, dtype=torch.float)
        actions_tensor[::self.num_dof] = actions.to(self.device).squeeze() * self.max_push_effort
        self.sim.set_indexed("set_dof_actuation_force_tensor", actions_tensor, None, None)

    def post_physics_step(self):
        """cartpole.py:165-173 as: device-side nonzero() + ONE fused kernel."""
        stream = self._stream()
        if self.rand_mode != "philox":
            # parity mode: the uniforms are drawn on the host side in the reference's order, which needs
            # len(env_ids) first -- the same host sync the reference has
            self.compact_resets()
            k = int(self.reset_count.item())          # same host sync as `len(env_ids)` cartpole.py:169
            if k > 0 and self._u0 is None:
                self._u0, self._u1 = self.draw_reset_uniforms(k)
        p, b = self.abi_structs()
        _lib.call("ige_cartpole_post_physics", p, b, stream=stream)
        self.sim.set_indexed("set_dof_state_tensor_indexed", self.dof_state, self.reset_env_ids, self.reset_count)
        self._u0 = self._u1 = None
