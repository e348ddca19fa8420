"""Agent-side rollout bookkeeping of ``A2CAgentMA.play_steps``.

Mirror of the state ``play_steps`` keeps between environment steps
(/root/reference/isaacgymenvs/learning/A2CAgent_MA.py:35-56, same attribute names): ``current_rewards``,
``current_shaped_rewards``, ``current_lengths``, ``dones``, the experience-buffer rows ``rewards`` /
``dones`` and the three ``AverageMeter`` s ``game_rewards`` / ``game_shaped_rewards`` / ``game_lengths``.
``after_env_step`` is ONE C-ABI launch (``ige_ma_rollout_step``) and never synchronises: the number of
finished environments stays on the device (``done_count``); ``discount_values`` is the GAE pass of
:58-67 (``ige_ma_discount_values``).  No CPU fallback: CPU tensors raise ``IgeError``.
"""
import math

import torch

from . import _abi, _lib


class _Meter:
    """View of one device-resident AverageMeter (rl_games algos_torch/torch_ext.py)."""

    def __init__(self, owner, slot):
        self._o, self._s = owner, slot

    @property
    def mean(self):
        return self._o.meter_mean[self._s:self._s + 1]

    @property
    def current_size(self):
        return int(self._o.meter_size[self._s])  # host read: logging only

    def get_mean(self):
        return self.mean.cpu().numpy()


class RolloutBookkeeperMA:
    def __init__(self, num_rows, num_agents, horizon_length, device, gamma=0.99, tau=0.95, value_bootstrap=True,
                 games_to_track=100, scale_value=1.0, shift_value=0.0, min_val=-math.inf, max_val=math.inf):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.IgeError("RolloutBookkeeperMA needs a CUDA device: there is no CPU fallback")
        self.device, self.num_rows, self.num_agents, self.horizon_length = dev, num_rows, num_agents, horizon_length
        self.gamma, self.tau, self.value_bootstrap, self.games_to_track = gamma, tau, value_bootstrap, games_to_track
        self.scale_value, self.shift_value, self.min_val, self.max_val = scale_value, shift_value, min_val, max_val
        f = dict(device=dev, dtype=torch.float32)
        self.current_rewards = torch.zeros((num_rows, 1), **f)
        self.current_shaped_rewards = torch.zeros((num_rows, 1), **f)
        self.current_lengths = torch.zeros((num_rows,), **f)
        # experience buffer rows; 'dones' has one extra row: row n+1 is written by step n, row 0 carries over
        self.mb_rewards = torch.zeros((horizon_length, num_rows, 1), **f)
        self._mb_dones = torch.ones((horizon_length + 1, num_rows), device=dev, dtype=torch.uint8)
        self.mb_returns = torch.zeros((horizon_length, num_rows, 1), **f)
        self.mb_advs = torch.zeros((horizon_length, num_rows, 1), **f)
        self.done_rows = torch.zeros((max(num_rows, 1),), device=dev, dtype=torch.int32)
        self.done_count = torch.zeros((1,), device=dev, dtype=torch.int32)
        self.meter_mean = torch.zeros((4,), **f)
        self.meter_size = torch.zeros((4,), device=dev, dtype=torch.int32)
        self.game_rewards, self.game_shaped_rewards, self.game_lengths = (_Meter(self, i) for i in range(3))
        self._reduce = torch.zeros((4,), device=dev, dtype=torch.float64)
        words = _lib.lib().ige_rollout_scratch_words(max(num_rows, 1))
        self._scratch = torch.zeros((words,), device=dev, dtype=torch.int32)
        self._n = 0

    # -- views with the reference's names --------------------------------------------------------------------
    @property
    def dones(self):
        """``self.dones`` (uint8): the flags of the latest env step."""
        return self._mb_dones[self._n]

    @property
    def mb_dones(self):
        """experience_buffer.tensor_dict['dones'] (horizon, rows)."""
        return self._mb_dones[:self.horizon_length]

    def env_done_indices(self):
        """``all_done_indices[::num_agents]`` of the latest step, shape (k, 1).  Host sync (reads the count)."""
        k = int(self.done_count)
        return self.done_rows[:k].long().unsqueeze(1)

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def begin_rollout(self):
        """Start of play_steps: the last step's dones become row 0 of the next buffer."""
        if self._n:
            self._mb_dones[0].copy_(self._mb_dones[self._n])
        self._n = 0

    def after_env_step(self, rewards, dones, time_outs=None, values=None):
        """Lines :34-56 for step ``n`` (the call order within play_steps is kept by the caller)."""
        n = self._n
        if n >= self.horizon_length:
            raise _lib.IgeError("after_env_step called more than horizon_length times; call begin_rollout()")
        boot = bool(self.value_bootstrap and time_outs is not None)
        a = _abi.IgeRolloutStepArgs()
        a.num_rows, a.num_agents, a.value_bootstrap = self.num_rows, self.num_agents, int(boot)
        a.meter_max_size, a.gamma = self.games_to_track, self.gamma
        a.shaper_shift, a.shaper_scale = self.shift_value, self.scale_value
        a.shaper_min, a.shaper_max = self.min_val, self.max_val
        rewards = rewards.reshape(-1)
        for name, t, dt in (("rewards", rewards, torch.float32), ("dones", dones, torch.int64)):
            if t.device != self.device or t.dtype != dt or not t.is_contiguous() or t.numel() != self.num_rows:
                raise _lib.IgeError(f"{name}: expected a contiguous {dt} tensor of {self.num_rows} rows on {self.device}")
        a.rewards, a.dones = rewards.data_ptr(), dones.data_ptr()
        if boot:
            if time_outs.dtype == torch.bool:
                time_outs = time_outs.view(torch.uint8)
            elif time_outs.dtype != torch.uint8:
                time_outs = time_outs.to(torch.uint8)
            values, time_outs = values.reshape(-1), time_outs.reshape(-1)
            for name, t, dt in (("values", values, torch.float32), ("time_outs", time_outs, torch.uint8)):
                if t.device != self.device or t.dtype != dt or not t.is_contiguous() or t.numel() != self.num_rows:
                    raise _lib.IgeError(f"{name}: expected a contiguous {dt} tensor of {self.num_rows} rows on {self.device}")
            a.time_outs, a.values = time_outs.data_ptr(), values.data_ptr()
        a.mb_rewards = self.mb_rewards[n].data_ptr()
        a.dones_out = self._mb_dones[n + 1].data_ptr()
        a.current_rewards = self.current_rewards.data_ptr()
        a.current_shaped_rewards = self.current_shaped_rewards.data_ptr()
        a.current_lengths = self.current_lengths.data_ptr()
        a.done_rows, a.done_count = self.done_rows.data_ptr(), self.done_count.data_ptr()
        a.meter_mean, a.meter_size = self.meter_mean.data_ptr(), self.meter_size.data_ptr()
        a.reduce_scratch, a.scratch = self._reduce.data_ptr(), self._scratch.data_ptr()
        _lib.call("ige_ma_rollout_step", a, stream=self._stream())
        self._n = n + 1

    def discount_values(self, last_values, mb_values):
        """:58-67: mb_advs = GAE(dones, last_values, mb_dones, mb_values, mb_rewards); returns mb_returns."""
        a = _abi.IgeDiscountArgs()
        a.num_rows, a.horizon = self.num_rows, self.horizon_length
        a.gamma, a.gamma_tau = self.gamma, self.gamma * self.tau
        last_values, mb_values = last_values.reshape(-1), mb_values.reshape(self.horizon_length, self.num_rows)
        for t in (last_values, mb_values):
            if t.device != self.device or t.dtype != torch.float32 or not t.is_contiguous():
                raise _lib.IgeError("discount_values: expected contiguous float32 CUDA tensors")
        a.dones = self._mb_dones[self._n].data_ptr()
        a.last_values, a.mb_dones, a.mb_values = last_values.data_ptr(), self._mb_dones.data_ptr(), mb_values.data_ptr()
        a.mb_rewards, a.mb_advs, a.mb_returns = self.mb_rewards.data_ptr(), self.mb_advs.data_ptr(), self.mb_returns.data_ptr()
        _lib.call("ige_ma_discount_values", a, stream=self._stream())
        return self.mb_returns
