"""Drop-in binding of the C ABI under the REFERENCE's own task classes (executable).

The reference's plug-in point for the hot path is the Python virtual ``VecTask.post_physics_step``
(isaacgymenvs/tasks/base/vec_task.py:358-360, called from ``VecTask.step`` at :391).  ``bind(RefAnt)`` returns
``class FusedAnt(FusedLocomotionMixin, RefAnt)``: the reference class with exactly that one method replaced by a call
into ``libige_b200.so``.  Everything else -- ``__init__``, ``pre_physics_step``, ``VecTask.step``, the gymtorch tensor
views, ``obs_buf / rew_buf / reset_buf / progress_buf`` -- is the reference's own code and memory, untouched.

This module never imports the reference: the caller passes the class in (tests/test_gpu_binding.py does so with the
unmodified ``isaacgymenvs.tasks.ant.Ant``).
"""
import ctypes as C

import torch

from . import _abi, _lib


class FusedLocomotionMixin:
    """``class Ant(FusedLocomotionMixin, VecTask)`` / ``class Humanoid(FusedLocomotionMixin, VecTask)``:
    replaces ant.py:287-297 + :252-279 + :325-408 (humanoid.py:287-296 + :253-279 + :323-413).

    ``ige_rand_mode``: "philox" (production: reset noise drawn in the kernel, no host sync before the launch) or
    "injected" (the uniforms are drawn with ``torch.rand`` in the reference's call order and shapes -- what the
    parity test uses; needs ``len(env_ids)`` on the host exactly like the reference, ant.py:292)."""

    ige_rand_mode = "philox"
    _IGE_ENTRY = {8: ("ige_ant_post_physics", 24), 21: ("ige_humanoid_post_physics", 12)}

    def ige_setup(self):
        """Call once at the end of the reference ``__init__`` (all tensors exist)."""
        n, dev = self.num_envs, torch.device(self.device)
        if dev.type != "cuda":
            raise _lib.IgeError("the fused post_physics_step needs sim_device=cuda tensors; there is no CPU path")
        self._ige_entry, sensors = self._IGE_ENTRY[int(self.num_dof)]
        # The gym API takes gymtorch-unwrapped tensors, never torch tensors (ant.py:265-271).  isaacgym is already
        # imported by the reference module that defines the class this mixin is bound to.
        import importlib
        self._ige_unwrap = importlib.import_module("isaacgym.gymtorch").unwrap_tensor
        lib = _lib.lib()
        i32 = dict(dtype=torch.int32, device=dev)
        self.reset_env_ids = torch.zeros(n, **i32)
        self.reset_count = torch.zeros(1, **i32)
        self.reset_rank = torch.zeros(n, **i32)
        self.rng_step = torch.zeros(1, dtype=torch.int64, device=dev)
        self._ige_list_scratch = torch.zeros(lib.ige_reset_list_scratch_words(n), **i32)
        self._ige_compact_scratch = torch.zeros(lib.ige_compact_scratch_words(n), **i32)
        p = self._ige_p = _abi.IgeLocomotionParams()
        p.num_envs, p.num_dof, p.num_sensor_floats = n, int(self.num_dof), sensors
        p.dt, p.dof_vel_scale, p.contact_force_scale = self.dt, self.dof_vel_scale, self.contact_force_scale
        p.angular_velocity_scale = float(getattr(self, "angular_velocity_scale", 1.0))
        p.up_weight, p.heading_weight = self.up_weight, self.heading_weight
        p.actions_cost_scale, p.energy_cost_scale = self.actions_cost_scale, self.energy_cost_scale
        p.joints_at_limit_cost_scale = self.joints_at_limit_cost_scale
        p.termination_height, p.death_cost = self.termination_height, self.death_cost
        p.max_episode_length, p.clip_obs = float(self.max_episode_length), float(self.clip_obs)
        for name in ("targets", "inv_start_rot", "basis_vec0", "basis_vec1"):      # per-env copies of one row
            t = getattr(self, name)
            if not bool((t == t[0:1]).all()):
                raise ValueError(f"{name} differs between envs; the fused kernel takes it as a per-task constant")
            _abi.fill_array(getattr(p, name), t[0].tolist())
        _abi.fill_array(p.dof_limits_lower, self.dof_limits_lower.tolist())
        _abi.fill_array(p.dof_limits_upper, self.dof_limits_upper.tolist())
        if hasattr(self, "motor_efforts"):                                        # humanoid.py:355
            _abi.fill_array(p.motor_effort_ratio, (self.motor_efforts.detach().cpu() / self.max_motor_effort).tolist())
        b = self._ige_b = _abi.IgeLocomotionBuffers()
        for name in ("root_states", "dof_state", "vec_sensor_tensor", "initial_root_states", "initial_dof_pos",
                     "potentials", "prev_potentials", "up_vec", "heading_vec", "obs_buf", "rew_buf", "progress_buf",
                     "randomize_buf"):
            t = getattr(self, name)
            if not t.is_contiguous():
                raise _lib.IgeError(f"{name} must be contiguous")
            setattr(b, name, t.data_ptr())        # gymtorch views: borrowed for the call, never freed
        if hasattr(self, "dof_force_tensor"):
            b.dof_force_tensor = self.dof_force_tensor.data_ptr()
        b.list.ids, b.list.count = self.reset_env_ids.data_ptr(), self.reset_count.data_ptr()
        b.list.scratch = self._ige_list_scratch.data_ptr()
        return self

    def ige_draw_reset_uniforms(self, k):
        """injected mode: the two ``torch_rand_float`` draws of reset_idx (ant.py:257-258), same order and shapes."""
        return (torch.rand((k, self.num_dof), device=self.device), torch.rand((k, self.num_dof), device=self.device))

    def post_physics_step(self):
        self.gym.refresh_dof_state_tensor(self.sim)               # ant.py:233-235, unchanged
        self.gym.refresh_actor_root_state_tensor(self.sim)
        self.gym.refresh_force_sensor_tensor(self.sim)
        if hasattr(self, "dof_force_tensor"):
            self.gym.refresh_dof_force_tensor(self.sim)           # humanoid.py:292
        p, b = self._ige_p, self._ige_b
        b.actions = self.actions.data_ptr()                       # rebound every step (ant.py:282)
        b.reset_buf = self.reset_buf.data_ptr()                   # Humanoid rebinds self.reset_buf (humanoid.py:220)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        keep = None
        if self.ige_rand_mode == "philox":
            p.rand.mode, p.rand.seed, p.rand.step = _abi.IGE_RAND_PHILOX, int((getattr(self, "cfg", None) or {}).get("seed", 0)), self.rng_step.data_ptr()
        else:
            a = _abi.IgeCompactArgs()
            a.num_rows, a.group = self.num_envs, 1
            a.reset_buf, a.reset_ids = self.reset_buf.data_ptr(), self.reset_env_ids.data_ptr()
            a.reset_rank, a.count = self.reset_rank.data_ptr(), self.reset_count.data_ptr()
            a.scratch = self._ige_compact_scratch.data_ptr()
            _lib.call("ige_compact_resets", a, stream=stream)
            k = int(self.reset_count.item())                      # the reference's own host sync, len(env_ids) ant.py:292
            p.rand.mode, p.rand.reset_rank = _abi.IGE_RAND_INJECTED, self.reset_rank.data_ptr()
            if k > 0:
                keep = self.ige_draw_reset_uniforms(k)
                p.rand.u0, p.rand.u1 = keep[0].data_ptr(), keep[1].data_ptr()
            else:
                p.rand.u0 = p.rand.u1 = None
        _lib.call(self._ige_entry, p, b, stream=stream)
        self._ige_keep = keep                                     # alive until the stream has consumed them
        k = int(self.reset_count.item())                          # ant.py:292 (the gym API needs a host integer)
        if k > 0:                                                 # ant.py:263-271: ids are int32, ascending
            unwrap = self._ige_unwrap
            ids = unwrap(self.reset_env_ids[:k])
            self.gym.set_actor_root_state_tensor_indexed(self.sim, unwrap(self.initial_root_states), ids, k)
            self.gym.set_dof_state_tensor_indexed(self.sim, unwrap(self.dof_state), ids, k)
        if hasattr(self, "compute_true_objective") and int(self.num_dof) == 8:
            self.extras["true_objective"] = self.root_states[:, 7]   # ant.py:245-250


def bind(ref_cls, mixin=FusedLocomotionMixin, name=None):
    """``bind(isaacgymenvs.tasks.ant.Ant)`` -> the reference class with the fused ``post_physics_step``."""
    return type(name or "Fused" + ref_cls.__name__, (mixin, ref_cls), {})
