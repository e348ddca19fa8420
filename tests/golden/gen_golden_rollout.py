"""Golden fixtures for the agent-side rollout bookkeeping.

Runs the UNMODIFIED reference ``A2CAgentMA.play_steps`` (/root/reference/isaacgymenvs/learning/
A2CAgent_MA.py:10-76) on a scripted environment and records what it wrote.  ``rl_games`` (setup.py:23,
"rl-games>=1.6.0") is not installed in the container and not vendored by the reference, so the pieces of
it that play_steps calls are restated below from rl_games' published source (1.6.x):
``DefaultRewardsShaper.__call__`` (common/tr_helpers.py), ``AverageMeter`` (algos_torch/torch_ext.py),
``ExperienceBuffer.update_data`` (common/experience.py), ``A2CBase.env_step`` / ``discount_values`` /
``swap_and_flatten01`` (common/a2c_common.py).  Lines 35-56 of play_steps itself -- the path under test --
are the reference's own code, executed as is.

Container-only (needs /root/reference):  python tests/golden/gen_golden_rollout.py
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_ROOT = os.environ.get("IGE_REFERENCE_ROOT", "/root/reference")


# ---- restated rl_games pieces ---------------------------------------------------------------------------------
class DefaultRewardsShaper:
    def __init__(self, scale_value=1, shift_value=0, min_val=-np.inf, max_val=np.inf):
        self.scale_value, self.shift_value, self.min_val, self.max_val = scale_value, shift_value, min_val, max_val

    def __call__(self, reward):
        reward = reward + self.shift_value
        reward = reward * self.scale_value
        return torch.clamp(reward, self.min_val, self.max_val)


class AverageMeter:
    def __init__(self, in_shape, max_size):
        self.max_size = max_size
        self.current_size = 0
        self.mean = torch.zeros(in_shape, dtype=torch.float32)

    def update(self, values):
        size = values.size()[0]
        if size == 0:
            return
        new_mean = torch.mean(values.float(), dim=0)
        size = np.clip(size, 0, self.max_size)
        old_size = min(self.max_size - size, self.current_size)
        size_sum = old_size + size
        self.current_size = size_sum
        self.mean = (self.mean * old_size + new_mean * size) / size_sum


class ExperienceBuffer:
    def __init__(self, horizon, rows):
        self.tensor_dict = {
            "obses": torch.zeros((horizon, rows, 1)),
            "dones": torch.zeros((horizon, rows), dtype=torch.uint8),
            "rewards": torch.zeros((horizon, rows, 1)),
            "values": torch.zeros((horizon, rows, 1)),
            "actions": torch.zeros((horizon, rows, 1)),
        }

    def update_data(self, name, index, val):
        self.tensor_dict[name][index, :] = val

    def get_transformed_list(self, fn, tensor_list):
        return {k: fn(self.tensor_dict[k]) for k in tensor_list}


def swap_and_flatten01(arr):
    s = arr.size()
    return arr.transpose(0, 1).reshape(s[0] * s[1], *s[2:])


class A2CAgentStub:
    """The slice of rl_games' A2CBase that play_steps touches."""

    def env_step(self, actions):
        obs, rewards, dones, infos = self.vec_env.step(actions)
        return {"obs": obs}, rewards.unsqueeze(1), dones, infos  # value_size == 1

    def get_action_values(self, obs):
        n = self.vec_env.n
        return {"actions": torch.zeros((self.rows, 1)), "values": self.script["values"][n].unsqueeze(1)}

    def get_values(self, obs):
        return self.script["last_values"].unsqueeze(1)

    def cast_obs(self, obs):
        return obs

    def discount_values(self, fdones, last_extrinsic_values, mb_fdones, mb_extrinsic_values, mb_rewards):
        lastgaelam = 0
        mb_advs = torch.zeros_like(mb_rewards)
        for t in reversed(range(self.horizon_length)):
            if t == self.horizon_length - 1:
                nextnonterminal = 1.0 - fdones
                nextvalues = last_extrinsic_values
            else:
                nextnonterminal = 1.0 - mb_fdones[t + 1]
                nextvalues = mb_extrinsic_values[t + 1]
            nextnonterminal = nextnonterminal.unsqueeze(1)
            delta = mb_rewards[t] + self.gamma * nextvalues * nextnonterminal - mb_extrinsic_values[t]
            mb_advs[t] = lastgaelam = delta + self.gamma * self.tau * nextnonterminal * lastgaelam
        return mb_advs


def load_reference_agent():
    """Import learning/A2CAgent_MA.py unmodified with the stub base class in place of rl_games."""
    for name in ("rl_games", "rl_games.algos_torch", "rl_games.common"):
        sys.modules.setdefault(name, types.ModuleType(name))
    cont = types.ModuleType("rl_games.algos_torch.a2c_continuous")
    cont.A2CAgent = A2CAgentStub
    common = types.ModuleType("rl_games.common.a2c_common")
    common.swap_and_flatten01 = swap_and_flatten01
    sys.modules["rl_games.algos_torch.a2c_continuous"] = cont
    sys.modules["rl_games.common.a2c_common"] = common
    sys.modules["rl_games.algos_torch"].a2c_continuous = cont
    path = os.path.join(REFERENCE_ROOT, "isaacgymenvs", "learning", "A2CAgent_MA.py")
    spec = importlib.util.spec_from_file_location("ref_A2CAgent_MA", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.A2CAgentMA


# ---- scripted environment -------------------------------------------------------------------------------------
class ScriptedEnv:
    def __init__(self, script):
        self.script, self.n = script, 0

    def step(self, actions):
        s, n = self.script, self.n
        self.n += 1
        return torch.zeros((s["rewards"].shape[1], 1)), s["rewards"][n].clone(), s["dones"][n].clone(), \
            {"time_outs": s["time_outs"][n].clone()}


class Observer:
    def __init__(self, agent):
        self.agent, self.done_rows, self.trace = agent, [], []

    def process_infos(self, infos, env_done_indices):
        a = self.agent
        self.done_rows.append(env_done_indices.flatten().clone())
        self.trace.append((
            [float(m.mean.flatten()[0]) for m in (a.game_rewards, a.game_shaped_rewards, a.game_lengths)],
            [int(m.current_size) for m in (a.game_rewards, a.game_shaped_rewards, a.game_lengths)]))


def make_script(num_envs, A, H, seed, done_prob, paired, timeout_prob=0.3):
    g = torch.Generator().manual_seed(seed)
    rows = num_envs * A
    rewards = torch.randn((H, rows), generator=g) * 3.0
    values = torch.randn((H, rows), generator=g)
    if paired:  # all agents of an env flip together (what the MA tasks produce)
        d = (torch.rand((H, num_envs), generator=g) < done_prob).repeat_interleave(A, dim=1)
    else:       # ragged: the [::A] stride then runs over the LIST of done rows
        d = torch.rand((H, rows), generator=g) < done_prob
    dones = d.to(torch.long)
    time_outs = (d & (torch.rand((H, rows), generator=g) < timeout_prob)).to(torch.uint8)
    return {
        "rewards": rewards, "values": values, "dones": dones, "time_outs": time_outs,
        "last_values": torch.randn((rows,), generator=g),
        "init_current_rewards": torch.randn((rows,), generator=g) * 5.0,
        "init_current_shaped_rewards": torch.randn((rows,), generator=g) * 0.05,
        "init_current_lengths": torch.randint(0, 50, (rows,), generator=g).float(),
        "init_dones": (torch.rand((rows,), generator=g) < 0.1).to(torch.uint8),
    }


def run_case(cls, num_envs, A, H, seed, done_prob, paired, value_bootstrap, games_to_track,
             scale=0.01, shift=0.0, min_val=-np.inf, max_val=np.inf, gamma=0.99, tau=0.95):
    rows = num_envs * A
    script = make_script(num_envs, A, H, seed, done_prob, paired)
    ag = cls.__new__(cls)
    ag.script, ag.rows = script, rows
    ag.update_list = ["actions", "values"]
    ag.horizon_length, ag.num_agents = H, A
    ag.use_action_masks, ag.has_central_value = False, False
    ag.value_bootstrap, ag.gamma, ag.tau = value_bootstrap, gamma, tau
    ag.vec_env = ScriptedEnv(script)
    ag.obs = {"obs": torch.zeros((rows, 1))}
    ag.dones = script["init_dones"].clone()
    ag.experience_buffer = ExperienceBuffer(H, rows)
    ag.rewards_shaper = DefaultRewardsShaper(scale_value=scale, shift_value=shift, min_val=min_val, max_val=max_val)
    ag.current_rewards = script["init_current_rewards"].clone().unsqueeze(1)
    ag.current_shaped_rewards = script["init_current_shaped_rewards"].clone().unsqueeze(1)
    ag.current_lengths = script["init_current_lengths"].clone()
    ag.game_rewards = AverageMeter(1, games_to_track)
    ag.game_shaped_rewards = AverageMeter(1, games_to_track)
    ag.game_lengths = AverageMeter(1, games_to_track)
    ag.algo_observer = Observer(ag)
    ag.tensor_list = ["actions", "values"]
    ag.batch_size = H * rows
    batch = ag.play_steps()  # <- the reference's own code

    obs = ag.algo_observer
    done_rows = np.full((H, rows), -1, np.int32)
    done_count = np.zeros((H,), np.int32)
    for t, r in enumerate(obs.done_rows):
        done_count[t] = r.numel()
        done_rows[t, :r.numel()] = r.numpy()
    blob = {k: v.numpy() for k, v in script.items()}
    blob.update({
        "cfg_num_agents": A, "cfg_horizon": H, "cfg_value_bootstrap": int(value_bootstrap),
        "cfg_games_to_track": games_to_track, "cfg_gamma": gamma, "cfg_tau": tau,
        "cfg_scale": scale, "cfg_shift": shift, "cfg_min": min_val, "cfg_max": max_val,
        "out_mb_rewards": ag.experience_buffer.tensor_dict["rewards"].squeeze(-1).numpy(),
        "out_mb_dones": ag.experience_buffer.tensor_dict["dones"].numpy(),
        "out_current_rewards": ag.current_rewards.squeeze(1).numpy(),
        "out_current_shaped_rewards": ag.current_shaped_rewards.squeeze(1).numpy(),
        "out_current_lengths": ag.current_lengths.numpy(),
        "out_done_rows": done_rows, "out_done_count": done_count,
        "out_meter_mean": np.array([t[0] for t in obs.trace], np.float32),
        "out_meter_size": np.array([t[1] for t in obs.trace], np.int32),
        # (rows*H, 1) in swap_and_flatten01 order -> back to (H, rows)
        "out_mb_returns": batch["returns"].reshape(rows, H).t().contiguous().numpy(),
    })
    return blob


CASES = [
    dict(name="rollout_ma_n64x2_s0", num_envs=64, A=2, H=8, seed=0, done_prob=0.1, paired=True,
         value_bootstrap=True, games_to_track=100),
    dict(name="rollout_ma_n40x3_s1_ragged", num_envs=40, A=3, H=6, seed=1, done_prob=0.3, paired=False,
         value_bootstrap=True, games_to_track=7),
    dict(name="rollout_ma_n50x2_s2_noboot_clamp", num_envs=50, A=2, H=5, seed=2, done_prob=0.05, paired=True,
         value_bootstrap=False, games_to_track=100, scale=0.5, shift=0.25, min_val=-1.0, max_val=1.0),
    dict(name="rollout_sa_n1100x1_s3", num_envs=1100, A=1, H=4, seed=3, done_prob=0.02, paired=True,
         value_bootstrap=True, games_to_track=100),
]


def main():
    cls = load_reference_agent()
    for c in CASES:
        c = dict(c)
        name = c.pop("name")
        blob = run_case(cls, **c)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **blob)
        print(f"{name}: {os.path.getsize(path)} bytes, done rows/step {blob['out_done_count'].tolist()}")


if __name__ == "__main__":
    main()
