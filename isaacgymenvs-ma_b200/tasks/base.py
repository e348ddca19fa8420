"""VecTask mirror: buffers + step() contract of isaacgymenvs/tasks/base/vec_task.py, with the
post-physics work delegated to the fused CUDA kernels through the C ABI.

What is kept from the reference (so callers cannot tell the difference):
  * ``allocate_buffers`` shapes / dtypes (vec_task.py:302-326): obs_buf f32 (N, num_obs),
    rew_buf f32, reset_buf i64 = 1, timeout_buf, progress_buf i64, randomize_buf i64, extras.
  * ``step(actions) -> (obs_dict, rew, reset, extras)`` order of operations (vec_task.py:362-410).
  * ``reset()`` / ``reset_done()`` / ``reset_idx(env_ids)`` (vec_task.py:422-457; resets are deferred to the next
    fused launch, see ``reset_idx``).
What is different: the simulator.  Isaac Gym's PhysX binary is not available, so ``self.sim`` is a
``SyntheticSim`` that owns the gymtorch-layout tensors and *records* the ``set_*_indexed`` calls
(the same role ``gym`` plays in the reference).  A real integration binds the same
``post_physics_step`` onto the reference classes -- see INTEGRATION.md.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np
import torch

from .. import _abi, _lib


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class SyntheticSim:
    """Owns the simulator-side tensors (gymtorch layouts) and records indexed set calls.

    ``tensors`` maps the reference attribute names (root_states, dof_state, ...) to torch tensors
    on the sim device.  ``set_indexed(name, ids, count)`` plays the role of
    ``gym.set_*_tensor_indexed(sim, tensor, ids_int32, count)``: recorded, never applied
    (SURVEY.md section 8c FakeGym semantics)."""

    def __init__(self, tensors):
        self.tensors = tensors
        self.calls = []
        self.record = True

    def set_indexed(self, name, tensor, ids, count):
        if self.record:
            self.calls.append((name, tensor, ids, count))

    def simulate(self):
        pass


class VecTaskB200:
    """Mirror of Env + VecTask (vec_task.py:67-133, 207-457) minus viewer / DR / sim creation."""

    # True for the task classes whose step kernels add into the episode-statistics accumulator (`stats` in their
    # C-ABI buffers struct: Cartpole, Ant / Humanoid, ShadowHand / AllegroHand, Franka*MA).
    ACCUMULATES_STATS = False

    num_agents = 1

    def __init__(self, cfg, sim: SyntheticSim, device, rl_device=None, rand_mode="philox", seed=0,
                 fuse_epilogue=True, accumulate_stats=True):
        self.cfg = cfg
        self.sim = sim
        self.gym = sim
        self.device = torch.device(device)
        self.rl_device = torch.device(rl_device) if rl_device is not None else self.device
        self.num_environments = int(cfg["num_envs"])
        self.num_observations = int(cfg["num_obs"])
        self.num_states = int(cfg.get("num_states", 0))
        self.num_actions = int(cfg["num_actions"])
        self.control_freq_inv = int(cfg.get("control_freq_inv", 1))
        self.clip_obs = float(cfg.get("clip_obs", math.inf))
        self.clip_actions = float(cfg.get("clip_actions", math.inf))
        self.control_steps = 0
        # vec_task.py:373-374, :399-400: the noise lambdas on actions / observations are part of step() and run as one
        # CUDA launch each, fused with the clamp that follows (dr_noise.NoiseLambda -> ige_dr_noise).  Everything else
        # in the randomisation config (sim_params, actor_params: vec_task.py:715-) re-parameterises the simulator and
        # is outside the hot path: refused rather than silently dropped.
        self.dr_randomizations = {}
        self._dr_spec = cfg.get("dr_randomizations") or {}
        extra = sorted(k for k in self._dr_spec if k not in ("frequency", "observations", "actions"))
        if cfg.get("randomize") and not self._dr_spec:
            raise NotImplementedError("`randomize` without `dr_randomizations`: only the observation / action noise "
                                      "of step() (vec_task.py:373-374, :399-400) is part of the fused path")
        if extra:
            raise NotImplementedError(f"domain randomisation of {extra} (vec_task.py:715-) is simulator set-up, not part "
                                      "of the fused path; only `observations` / `actions` noise is supported")
        self.rand_mode = rand_mode
        self.seed = int(seed)
        self.fuse_epilogue = fuse_epilogue
        self.emit_reset_list = os.environ.get("IGE_NO_LIST", "0") != "1"   # debug knob
        self.obs_dict = {}
        self.allocate_buffers()
        n_rows = self.num_envs * self.num_agents
        dev = self.device
        # outputs of the fused step() epilogue (vec_task.py:396,404)
        self.timeout_bool = torch.zeros(n_rows, dtype=torch.bool, device=dev)
        self.obs_clamped = torch.zeros_like(self.obs_buf) if fuse_epilogue else None
        # reset compaction outputs (what gym.set_*_indexed consumes)
        self.reset_env_ids = torch.zeros(self.num_envs, dtype=torch.int32, device=dev)
        self.reset_rank = torch.zeros(self.num_envs, dtype=torch.int32, device=dev)
        self.reset_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.rng_step = torch.zeros(1, dtype=torch.int64, device=dev)
        words = int(_lib.lib().ige_compact_scratch_words(self.num_envs * self.num_agents))
        self._compact_scratch = torch.zeros(words, dtype=torch.int32, device=dev)
        self._list_scratch = torch.zeros(_abi.reset_list_scratch_words(self.num_envs), dtype=torch.int32, device=dev)
        # only the tasks whose C ABI carries a `stats` pointer accumulate (ACCUMULATES_STATS); for the others the
        # accumulator is None rather than a buffer of zeros that would be all-reduced as if it meant something
        self.stats_rows = (torch.zeros(_abi.IGE_STATS_ROWS * _abi.IGE_STATS_STRIDE, dtype=torch.float32, device=dev)
                           if accumulate_stats and self.ACCUMULATES_STATS else None)
        self._u0 = self._u1 = None
        self.last_step, self.last_rand_step = 0, 0
        for tensor_id, name in enumerate(("actions", "observations")):
            if self._dr_spec.get(name):
                from ..dr_noise import NoiseLambda
                clip = self.clip_actions if name == "actions" else self.clip_obs
                self.dr_randomizations[name] = {"noise_lambda": NoiseLambda(
                    self._dr_spec[name], dev, clip=clip, tensor_id=tensor_id, seed=self.seed, step=self.rng_step,
                    rand_mode=rand_mode)}
        if self.dr_randomizations.get("observations") and self.obs_clamped is None:
            self.obs_clamped = torch.zeros_like(self.obs_buf)

    @property
    def stats(self):
        """Episode statistics accumulated by the step kernels: [IGE_NUM_STATS] column sums."""
        if self.stats_rows is None:
            return None
        return self.stats_rows.view(_abi.IGE_STATS_ROWS, _abi.IGE_STATS_STRIDE)[:, :_abi.IGE_NUM_STATS].sum(0)

    # -- vec_task.py:135-160 properties ------------------------------------------------------
    @property
    def num_envs(self):
        return self.num_environments

    @property
    def num_obs(self):
        return self.num_observations

    @property
    def num_acts(self):
        return self.num_actions

    # -- vec_task.py:302-326 -------------------------------------------------------------------
    def allocate_buffers(self):
        dev, n = self.device, self.num_envs * self.num_agents
        self.obs_buf = torch.zeros((n, self.num_obs), device=dev, dtype=torch.float)
        self.states_buf = torch.zeros((self.num_envs, self.num_states), device=dev, dtype=torch.float)
        self.rew_buf = torch.zeros(n, device=dev, dtype=torch.float)
        self.reset_buf = torch.ones(n, device=dev, dtype=torch.long)
        self.timeout_buf = torch.zeros(n, device=dev, dtype=torch.long)
        self.progress_buf = torch.zeros(n, device=dev, dtype=torch.long)
        self.randomize_buf = torch.zeros(self.num_envs, device=dev, dtype=torch.long)
        self.extras = {}

    # -- stream / launch helpers ---------------------------------------------------------------
    def _stream(self):
        if self.device.type != "cuda":
            raise _lib.IgeError(
                f"post_physics_step needs CUDA tensors (device={self.device}); there is no CPU path")
        return torch.cuda.current_stream(self.device).cuda_stream

    def _rand_struct(self):
        r = _abi.IgeResetRand()
        if self.rand_mode == "philox":
            r.mode = _abi.IGE_RAND_PHILOX
            r.seed = self.seed
            r.step = _ptr(self.rng_step)
        else:
            r.mode = _abi.IGE_RAND_INJECTED
            r.reset_rank = _ptr(self.reset_rank)
            r.u0 = _ptr(self._u0)
            r.u1 = _ptr(self._u1)
        return r

    def _list_struct(self):
        """Reset index list emitted by the fused kernel itself (ids ascending + device count)."""
        l = _abi.IgeResetList()
        if not self.emit_reset_list:
            return l
        l.ids = _ptr(self.reset_env_ids)
        l.count = _ptr(self.reset_count)
        l.scratch = _ptr(self._list_scratch)
        return l

    def compact_resets(self, group=1):
        """reset_buf.nonzero() on the device: fills reset_env_ids / reset_rank / reset_count."""
        a = _abi.IgeCompactArgs()
        a.num_rows = self.reset_buf.numel()
        a.group = group
        a.reset_buf = _ptr(self.reset_buf)
        a.reset_ids = _ptr(self.reset_env_ids)
        a.reset_rank = _ptr(self.reset_rank)
        a.count = _ptr(self.reset_count)
        a.scratch = _ptr(self._compact_scratch)
        _lib.call("ige_compact_resets", a, stream=self._stream())

    def draw_reset_uniforms(self, k):
        """INJECTED mode: the torch.rand calls reset_idx would make, in order.  Subclasses override."""
        raise NotImplementedError

    def set_injected_uniforms(self, u0, u1):
        self._u0, self._u1 = u0, u1

    # -- abstract task hooks -------------------------------------------------------------------
    def pre_physics_step(self, actions):
        raise NotImplementedError

    def post_physics_step(self):
        raise NotImplementedError

    def hot_step(self):
        """Exactly the hot path of one control step (what bench.py times): the reset + obs + reward +
        done work, nothing else.  Tasks that reset in pre-physics position override this."""
        self.post_physics_step()

    # -- vec_task.py:362-410 -------------------------------------------------------------------
    def step(self, actions):
        # vec_task.py:373; tasks whose pre-physics kernel clamps itself (FUSED_ACTION_CLAMP) skip the extra pass
        act_noise = self.dr_randomizations.get("actions")
        if act_noise:                                   # vec_task.py:373-376: noise, then the clamp, one launch
            if self.control_steps - self.last_rand_step >= self._dr_spec.get("frequency", 1) or self.control_steps == 0:
                self.last_rand_step = self.control_steps          # apply_randomizations re-derives the schedule
                for d in self.dr_randomizations.values():
                    d["noise_lambda"].set_schedule(self.control_steps)
            lam = act_noise["noise_lambda"]
            actions = actions.to(self.device, torch.float32).contiguous()
            clamped = torch.empty_like(actions)
            noisy = lam(actions, out_clamped=clamped, **self._dr_draws("actions", actions))
            action_tensor = noisy if self.FUSED_ACTION_CLAMP else clamped
        else:
            action_tensor = actions if self.FUSED_ACTION_CLAMP else torch.clamp(actions, -self.clip_actions, self.clip_actions)
        self.pre_physics_step(action_tensor)
        for _ in range(self.control_freq_inv):
            self.sim.simulate()
        self.post_physics_step()
        self.control_steps += 1
        obs_noise = self.dr_randomizations.get("observations")
        if obs_noise:                                   # vec_task.py:399-404: obs_buf <- noise(obs_buf); obs = clamp(obs_buf)
            if not act_noise and (self.control_steps - 1 - self.last_rand_step >= self._dr_spec.get("frequency", 1)
                                  or self.control_steps == 1):
                self.last_rand_step = self.control_steps - 1
                obs_noise["noise_lambda"].set_schedule(self.control_steps - 1)
            obs_noise["noise_lambda"](self.obs_buf, out=self.obs_buf, out_clamped=self.obs_clamped,
                                      **self._dr_draws("observations", self.obs_buf))
            self.timeout_buf = (self.timeout_bool if self.fuse_epilogue else
                                (self.progress_buf >= self.max_episode_length - 1) & (self.reset_buf != 0))
            obs = self.obs_clamped
        elif self.fuse_epilogue:
            # produced by the same kernel launch (obs_clamped / timeout_bool outputs)
            self.timeout_buf = self.timeout_bool
            obs = self.obs_clamped
        else:
            self.timeout_buf = (self.progress_buf >= self.max_episode_length - 1) & (self.reset_buf != 0)
            obs = torch.clamp(self.obs_buf, -self.clip_obs, self.clip_obs)
        self.extras["time_outs"] = self.timeout_buf.to(self.rl_device)
        self.obs_dict["obs"] = obs.to(self.rl_device)
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict, self.rew_buf.to(self.rl_device), self.reset_buf.to(self.rl_device), self.extras

    def _dr_draws(self, name, tensor):
        """injected parity mode: the draws the reference's closure would make, in its order -- on the first call
        ``corr = randn_like`` (vec_task.py:678-679), then ``randn_like`` / ``rand_like`` (:684 / :710)."""
        if self.rand_mode != "injected":
            return {}
        lam = self.dr_randomizations[name]["noise_lambda"]
        kw = {}
        if lam.corr is None:
            kw["corr"] = torch.randn(tensor.shape)
        kw["draw"] = torch.randn(tensor.shape) if lam.spec["distribution"] == "gaussian" else torch.rand(tensor.shape)
        return kw

    def get_state(self):
        return torch.clamp(self.states_buf, -self.clip_obs, self.clip_obs).to(self.rl_device)

    def zero_actions(self):
        return torch.zeros([self.num_envs, self.num_actions], dtype=torch.float32, device=self.rl_device)

    def reset(self):
        self.obs_dict["obs"] = torch.clamp(self.obs_buf, -self.clip_obs, self.clip_obs).to(self.rl_device)
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict

    def reset_idx(self, env_ids):
        """vec_task.py:422-426 virtual.  The fused kernels run ``reset_idx`` for every flagged env INSIDE the next
        ``post_physics_step`` (tasks that reset in pre-physics: inside ``pre_physics_step``), so resetting specific
        envs means flagging them: the reset -- state rows, noise, cleared flags -- then happens in the same launch
        as the observations, exactly where the reference's ``step()`` would run it."""
        ids = torch.as_tensor(env_ids, device=self.device).long().flatten()
        if self.num_agents > 1:      # agent-axis tasks: env e owns rows e*A .. e*A + A - 1 (franka_reach_MA.py:885-889)
            ids = (ids[:, None] * self.num_agents + torch.arange(self.num_agents, device=self.device)[None, :]).flatten()
        self.reset_buf[ids] = 1

    def reset_done(self):
        """vec_task.py:442-457: ``(obs_dict, done_env_ids)``; called by RLGPUEnv.reset_done
        (utils/rlgames_utils.py:252-253).  ``done_env_ids`` are the rows whose ``reset_buf`` is set, as in the
        reference; their reset is deferred to the next step's fused launch (see ``reset_idx``)."""
        done_env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        if len(done_env_ids) > 0 and self.num_agents == 1:
            self.reset_idx(done_env_ids)
        self.obs_dict["obs"] = torch.clamp(self.obs_buf, -self.clip_obs, self.clip_obs).to(self.rl_device)
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict, done_env_ids

    def get_number_of_agents(self):
        """What RLGPUEnv.get_number_of_agents forwards to (utils/rlgames_utils.py:255-256): vec_task.py:102 num_agents."""
        return self.num_agents

    # -- pipeline=cpu: simulator state lives in pinned host memory ------------------------------
    FUSED_ACTION_CLAMP = False
    HOST_SIM_ATTRS = ()      # task attributes that alias simulator / policy memory (set by subclasses)
    HOST_OUT_ATTRS = ("obs_buf", "rew_buf", "reset_buf")
    HOST_INOUT_ATTRS = ()    # subset of HOST_SIM_ATTRS that reset_idx rewrites in place

    def map_host_sim(self, extra=()):
        """``sim_device=cpu pipeline=cpu`` (vec_task.py:82-88) without a staging copy: the simulator
        tensors, the actions and the three step() results are re-homed in PINNED host memory and the
        fused kernel reads / writes them over PCIe itself (pinned allocations are device-addressable
        under UVA; the TMA / coalesced accesses of the kernel become the PCIe transactions, reads and
        writes overlap on the full-duplex link).  Task-owned state (progress, potentials, targets ...)
        stays in HBM.  Returns {name: pinned tensor}; the tensors stay ordinary CPU tensors for the host
        side.  Only the kernel launch touches them afterwards -- there is still no CPU compute path."""
        if self.device.type != "cuda":
            raise _lib.IgeError("map_host_sim() needs a CUDA device to run the kernels on")
        names = [n for n in tuple(self.HOST_SIM_ATTRS) + tuple(self.HOST_OUT_ATTRS) + tuple(extra)
                 if getattr(self, n, None) is not None]
        moved = self._rehome(names, lambda t: torch.empty(t.shape, dtype=t.dtype, pin_memory=True))
        self.host_mapped = moved
        return moved

    def _rehome(self, names, make):
        """Replace the tensors behind the attributes `names` (and their entries in sim.tensors) by
        ``make(old_tensor)`` holding the same values.  Returns {name: new tensor}."""
        moved = {}
        for name in names:
            t = getattr(self, name)
            h = make(t)
            h.copy_(t)
            setattr(self, name, h)
            for k, v in list(self.sim.tensors.items()):
                if v is t:
                    self.sim.tensors[k] = h
            moved[name] = h
        torch.cuda.synchronize(self.device)
        self._on_remap()
        return moved

    def _on_remap(self):
        pass

    # -- helpers for subclasses ----------------------------------------------------------------
    def _adopt(self, st, names):
        """Move numpy state arrays onto the device under the reference attribute names."""
        for k in names:
            v = st[k]
            t = torch.from_numpy(np.ascontiguousarray(v)).to(self.device) if isinstance(v, np.ndarray) else v
            setattr(self, k, t)

    @staticmethod
    def _row_constant(t, name):
        """The reference keeps per-env copies of constants (ant.py:103-114); the kernels take them as
        scalars, so refuse anything that is not actually constant over envs."""
        if t.dim() == 2 and t.shape[0] > 1 and not bool((t == t[0:1]).all()):
            raise ValueError(f"{name} differs between envs; the fused kernel takes it as a per-task constant")
        return [float(x) for x in (t[0] if t.dim() == 2 else t).tolist()]
