"""ShadowHand mirror (reference: isaacgymenvs/tasks/shadow_hand.py).

The reset section of ``pre_physics_step`` (:670-682) and the whole ``post_physics_step`` (:710-715)
are one C-ABI call each; the action -> position-target arithmetic of ``pre_physics_step``
(:684-698) is outside the hot path and stays plain torch.
"""
import torch

from .. import _abi, _lib
from .base import VecTaskB200, _ptr

NUM_OBS = {"openai": 42, "full_no_vel": 77, "full": 157, "full_state": 211}


class ShadowHand(VecTaskB200):
    ACCUMULATES_STATS = True
    FUSED_ACTION_CLAMP = True
    NUM_OBS = NUM_OBS
    NUM_STATES = 211
    OBS_IDS = _abi.IGE_HAND_OBS

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        self.obs_type = cfg["obs_type"]
        self.asymmetric_obs = bool(cfg["asymmetric_obs"])
        cfg["num_obs"] = self.NUM_OBS[self.obs_type]       # shadow_hand.py:109-133, allegro_hand.py:108-126
        cfg["num_states"] = self.NUM_STATES if self.asymmetric_obs else 0
        super().__init__(cfg, sim, device, **kw)
        for k in ("dist_reward_scale", "rot_reward_scale", "rot_eps", "action_penalty_scale", "reach_goal_bonus",
                  "fall_dist", "fall_penalty", "success_tolerance", "av_factor", "vel_obs_scale",
                  "force_torque_obs_scale", "reset_position_noise", "reset_dof_pos_noise", "reset_dof_vel_noise",
                  "dt", "act_moving_average"):
            setattr(self, k, float(cfg[k]))
        self.max_episode_length = int(cfg["max_episode_length"])
        self.max_consecutive_successes = int(cfg["max_consecutive_successes"])
        self.ignore_z = bool(cfg.get("ignore_z_rot", False))
        self.use_relative_control = bool(cfg["use_relative_control"])
        self.shadow_hand_dof_speed_scale = float(cfg["dof_speed_scale"])
        self.num_shadow_hand_dofs = int(cfg["num_hand_dofs"])
        self.num_fingertips = int(cfg["num_fingertips"])
        self.up_axis_idx = 2
        self.goal_displacement = [float(x) for x in cfg["goal_displacement"]]
        self.force_prob_range_f = [float(x) for x in cfg["force_prob_range"]]
        t, n = sim.tensors, self.num_envs
        self.dof_state = t["dof_state"]                                   # shadow_hand.py:167-170
        self.num_dofs = self.dof_state.shape[0] // n
        self.shadow_hand_dof_state = self.dof_state.view(n, -1, 2)[:, :self.num_shadow_hand_dofs]
        self.shadow_hand_dof_pos = self.shadow_hand_dof_state[..., 0]
        self.shadow_hand_dof_vel = self.shadow_hand_dof_state[..., 1]
        self.root_state_tensor = t["root_state_tensor"]
        self.rigid_body_states = t["rigid_body_states"]
        self.num_bodies = self.rigid_body_states.shape[1]
        self.vec_sensor_tensor = t["vec_sensor_tensor"]
        self.dof_force_tensor = t["dof_force_tensor"]
        dev = self.device
        self.actions = torch.zeros((n, self.num_actions), device=dev)
        self.reset_goal_buf = self.reset_buf.clone()                      # shadow_hand.py:190
        self.actor_dirty = torch.zeros(self.root_state_tensor.shape[0], dtype=torch.long, device=dev)
        self.goal_rank = torch.zeros(n, dtype=torch.int32, device=dev)
        self.goal_env_ids = torch.zeros(n, dtype=torch.int32, device=dev)
        self.goal_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.root_reset_ids = torch.zeros(self.root_state_tensor.shape[0], dtype=torch.int32, device=dev)
        self.root_reset_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self._compact_scratch_actors = torch.zeros(2 + (self.root_state_tensor.shape[0] + 2047) // 2048,
                                                   dtype=torch.int32, device=dev)
        self._reduce_scratch = torch.zeros(4, dtype=torch.int32, device=dev)
        self.states_clamped = torch.zeros_like(self.states_buf) if self.fuse_epilogue else None
        self._u = {}

    def adopt_state(self, st):
        self._adopt(st, ["goal_states", "goal_init_state", "object_init_state", "hand_indices", "object_indices",
                         "goal_object_indices", "fingertip_handles", "actuated_dof_indices",
                         "shadow_hand_dof_lower_limits", "shadow_hand_dof_upper_limits",
                         "shadow_hand_dof_default_pos", "shadow_hand_dof_default_vel", "prev_targets", "cur_targets",
                         "rb_forces", "random_force_prob", "successes", "consecutive_successes"])
        for k in ("reset_buf", "reset_goal_buf", "progress_buf", "randomize_buf", "actions"):
            getattr(self, k).copy_(torch.from_numpy(st[k]).to(self.device))
        self.shadow_hand_default_dof_pos = torch.zeros(self.num_shadow_hand_dofs, device=self.device)  # :166
        # host copies of the small constant tensors (read once: no D2H inside a step / graph capture)
        self._consts = dict(
            fingertips=[int(h) for h in self.fingertip_handles.tolist()],
            lower=self.shadow_hand_dof_lower_limits.tolist(), upper=self.shadow_hand_dof_upper_limits.tolist(),
            default_pos=self.shadow_hand_dof_default_pos.tolist(), default_vel=self.shadow_hand_dof_default_vel.tolist(),
            reset_base=self.shadow_hand_default_dof_pos.tolist())

    # ------------------------------------------------------------------------------------------
    def abi_structs(self):
        p = _abi.IgeShadowHandParams()
        p.num_envs = self.num_envs
        p.num_hand_dofs, p.num_dofs_per_env = self.num_shadow_hand_dofs, self.num_dofs
        p.num_bodies, p.num_actions = self.num_bodies, self.num_actions
        p.obs_type = self.OBS_IDS[self.obs_type]
        p.asymmetric_obs = int(self.asymmetric_obs)
        p.ignore_z_rot = int(self.ignore_z)
        p.max_consecutive_successes = self.max_consecutive_successes
        c = self._consts
        for i, h in enumerate(c["fingertips"]):
            p.fingertip_handles[i] = h
        p.max_episode_length = float(self.max_episode_length)
        for k in ("dist_reward_scale", "rot_reward_scale", "rot_eps", "action_penalty_scale", "success_tolerance",
                  "reach_goal_bonus", "fall_dist", "fall_penalty", "av_factor", "vel_obs_scale",
                  "force_torque_obs_scale", "reset_position_noise", "reset_dof_pos_noise", "reset_dof_vel_noise"):
            setattr(p, k, getattr(self, k))
        p.clip_obs = self.clip_obs
        _abi.fill_array(p.goal_displacement, self.goal_displacement)
        _abi.fill_array(p.force_prob_range, self.force_prob_range_f)
        _abi.fill_array(p.dof_lower, c["lower"])
        _abi.fill_array(p.dof_upper, c["upper"])
        _abi.fill_array(p.dof_default_pos, c["default_pos"])
        _abi.fill_array(p.dof_default_vel, c["default_vel"])
        _abi.fill_array(p.dof_reset_base, c["reset_base"])
        r = p.rand
        if self.rand_mode == "philox":
            r.mode, r.seed, r.step = _abi.IGE_RAND_PHILOX, self.seed, _ptr(self.rng_step)
        else:
            r.mode = _abi.IGE_RAND_INJECTED
            r.goal_rank, r.env_rank = _ptr(self.goal_rank), _ptr(self.reset_rank)
            for k in "ABCD":
                setattr(r, "u" + k, _ptr(self._u.get(k)))
        b = _abi.IgeShadowHandBuffers()
        for name in ("dof_state", "root_state_tensor", "rigid_body_states", "vec_sensor_tensor", "dof_force_tensor",
                     "actions", "hand_indices", "object_indices", "goal_object_indices", "goal_states",
                     "goal_init_state", "object_init_state", "prev_targets", "cur_targets", "rb_forces",
                     "random_force_prob", "successes", "consecutive_successes", "obs_buf", "rew_buf", "reset_buf",
                     "reset_goal_buf", "progress_buf", "randomize_buf", "actor_dirty"):
            setattr(b, name, _ptr(getattr(self, name)))
        b.states_buf = _ptr(self.states_buf) if self.asymmetric_obs else None
        b.obs_clamped = _ptr(self.obs_clamped) if self.fuse_epilogue else None
        b.states_clamped = _ptr(self.states_clamped) if (self.fuse_epilogue and self.asymmetric_obs) else None
        b.timeout_buf = _ptr(self.timeout_bool) if self.fuse_epilogue else None
        b.stats = _ptr(self.stats_rows)
        b.reduce_scratch = _ptr(self._reduce_scratch)
        b.list = self._list_struct()
        return p, b

    # ------------------------------------------------------------------------------------------
    def set_injected_uniforms(self, **u):
        self._u = {k: v for k, v in u.items() if v is not None}

    def draw_reset_uniforms(self, k_goal, k_env):
        """torch.rand calls of pre_physics_step in the reference's order (shadow_hand.py:587,610,613,644)."""
        u = {}
        if k_goal > 0:
            u["A"] = torch.rand((k_goal, 4), device=self.device)
        if k_env > 0:
            u["B"] = torch.rand((k_env, self.num_shadow_hand_dofs * 2 + 5), device=self.device)
            u["C"] = torch.rand((k_env, 4), device=self.device)
            u["D"] = torch.rand(k_env, device=self.device)
        return u

    def _compact(self, flags, ids, rank, count, scratch):
        a = _abi.IgeCompactArgs()
        a.num_rows, a.group = flags.numel(), 1
        a.reset_buf, a.reset_ids, a.reset_rank, a.count = _ptr(flags), _ptr(ids), _ptr(rank), _ptr(count)
        a.scratch = _ptr(scratch)
        _lib.call("ige_compact_resets", a, stream=self._stream())

    def reset_resolved(self):
        """shadow_hand.py:670-682 as one kernel (+ the device-side nonzero() passes the parity mode needs)."""
        stream = self._stream()
        if self.rand_mode != "philox":
            self.compact_resets()
            self._compact(self.reset_goal_buf, self.goal_env_ids, self.goal_rank, self.goal_count, self._compact_scratch)
            k_env, k_goal = int(self.reset_count.item()), int(self.goal_count.item())
            if not self._u and (k_env or k_goal):
                self._u = self.draw_reset_uniforms(k_goal, k_env)
        self.actor_dirty.zero_()
        p, b = self.abi_structs()
        _lib.call("ige_shadow_hand_reset", p, b, stream=stream)
        self._u = {}

    def sim_reset_calls(self):
        """The gym.set_*_indexed calls of reset_target_pose / reset_idx (:597-600, :634-664): ids come
        from compacting the dirty-actor flags (== sorted unique()) and hand_indices[env_ids]."""
        self._compact(self.actor_dirty, self.root_reset_ids, None, self.root_reset_count, self._compact_scratch_actors)
        self.sim.set_indexed("set_actor_root_state_tensor_indexed", self.root_state_tensor, self.root_reset_ids,
                             self.root_reset_count)
        hand_ids = self.hand_indices[self.reset_env_ids.long()].to(torch.int32)
        self.sim.set_indexed("set_dof_position_target_tensor_indexed", self.prev_targets, hand_ids, self.reset_count)
        self.sim.set_indexed("set_dof_state_tensor_indexed", self.dof_state, hand_ids, self.reset_count)

    def pre_physics_step(self, actions):
        self.reset_resolved()
        if self.sim.record:
            self.sim_reset_calls()
        self.apply_actions(actions)

    def target_args(self, actions):
        """IgeHandTargetArgs for `actions` (kept separate so that tests can hand the same struct to the oracle)."""
        src = actions.to(self.device, dtype=torch.float32).contiguous()
        if src.data_ptr() == self.actions.data_ptr():
            src = src.clone()
        a = _abi.IgeHandTargetArgs()
        a.num_envs, a.num_actions = self.num_envs, self.num_actions
        a.num_dofs_per_env = self.cur_targets.shape[1]
        a.use_relative_control = int(self.use_relative_control)
        a.clip_actions = self.clip_actions
        a.speed_scale_dt = self.shadow_hand_dof_speed_scale * self.dt      # the Python-float product, then fp32
        a.act_moving_average = self.act_moving_average
        a.one_minus_mavg = 1.0 - self.act_moving_average
        idx = self.actuated_dof_indices.tolist()
        _abi.fill_array(a.actuated, idx)
        _abi.fill_array(a.lower, self.shadow_hand_dof_lower_limits[self.actuated_dof_indices].tolist())
        _abi.fill_array(a.upper, self.shadow_hand_dof_upper_limits[self.actuated_dof_indices].tolist())
        a.actions_in, a.actions = _ptr(src), _ptr(self.actions)
        a.prev_targets, a.cur_targets = _ptr(self.prev_targets), _ptr(self.cur_targets)
        return a, src

    def apply_actions(self, actions):
        """shadow_hand.py:684-698 (+ the clamp of VecTask.step, vec_task.py:373) as one launch: actions ->
        position targets of the actuated DOFs (absolute + moving average, or relative), limit clamp,
        prev_targets update."""
        a, src = self.target_args(actions)
        _lib.call("ige_shadow_hand_pre_physics_targets", a, stream=self._stream())
        self._actions_in = src
        self.sim.set_indexed("set_dof_position_target_tensor", self.cur_targets, None, None)

    def hot_step(self):
        self.reset_resolved()          # pre-physics position (shadow_hand.py:670-682)
        self.post_physics_step()       # post-physics position (:710-715)

    def post_physics_step(self):
        """shadow_hand.py:710-715 as ONE fused kernel."""
        p, b = self.abi_structs()
        _lib.call("ige_shadow_hand_post_physics", p, b, stream=self._stream())
        # shadow_hand.py:424 `.mean()` of the 1-element tensor == the element: a 0-dim view, no extra kernel
        self.extras["consecutive_successes"] = self.consecutive_successes.reshape(())


class AllegroHand(ShadowHand):
    """AllegroHand mirror (reference: isaacgymenvs/tasks/allegro_hand.py) -- SURVEY.md 8f row 2.  ``reset_idx``
    (:507-573), ``reset_target_pose`` (:491-505), the action -> target arithmetic (:590-606) and ``compute_hand_reward``
    (:664-729) are ShadowHand's lines on a 16-DOF hand whose joints are all actuated; the observation layouts
    (:439-489: 50 / 72 / 88 wide, no fingertip or force-sensor segments) are three more instantiations of the same
    kernels (csrc/shadow_hand.cu, IGE_HAND_OBS_ALLEGRO_*)."""
    NUM_OBS = {"full_no_vel": 50, "full": 72, "full_state": 88}
    NUM_STATES = 88
    OBS_IDS = _abi.IGE_HAND_OBS_ALLEGRO
