"""Generate the committed golden fixtures from the UNMODIFIED reference (container-only).

    python tests/golden/gen_golden.py            # writes tests/golden/<task>_n<N>_s<seed>.npz

Each fixture is a short trajectory: synthetic state (package ``synthetic``) -> reference
``VecTask.step`` (tests/ref_harness.py) -> every buffer the path writes, plus the uniforms
``reset_idx`` drew and the int32 index lists it handed to ``gym.set_*_indexed``.
The GPU box has no /root/reference: tests there read only these .npz files.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import ref_harness  # noqa: E402
from ige_loader import load_package  # noqa: E402

pkg = load_package()
syn = pkg.synthetic

CFGS = {"cartpole": syn.CARTPOLE_CFG, "ant": syn.ANT_CFG, "humanoid": syn.HUMANOID_CFG,
        "shadow_hand": syn.SHADOW_HAND_CFG, "allegro_hand": syn.ALLEGRO_HAND_CFG}
for _v in syn.FRANKA_VARIANTS:
    CFGS["franka_" + _v] = None     # built per fixture: synthetic.franka_ma_cfg(variant, A, T)

# (task, num_envs, seed, reset_prob of step 0, steps)
FIXTURES = [
    ("cartpole", 512, 0, 0.02, 3),      # BASELINE.json configs[0]
    ("cartpole", 64, 1, 1.0, 2),
    ("ant", 128, 0, 0.05, 3),
    ("ant", 64, 1234, 1.0, 2),
    ("humanoid", 96, 0, 0.05, 3),
    ("humanoid", 64, 1234, 1.0, 2),
    ("shadow_hand", 64, 0, 0.05, 3, dict(obs_type="full_state")),
    ("shadow_hand", 48, 1, 1.0, 2, dict(obs_type="full_state")),            # first step: everything resets
    ("shadow_hand", 48, 2, 0.1, 2, dict(obs_type="openai", asymmetric_obs=True)),
    ("shadow_hand", 48, 3, 0.1, 2, dict(obs_type="full_no_vel")),
    ("shadow_hand", 48, 4, 0.1, 2, dict(obs_type="full", max_consecutive_successes=3)),
    ("shadow_hand", 48, 5, 0.0, 2, dict(obs_type="full_state", goal_reset_prob=0.2)),   # goal-only resets
    ("allegro_hand", 48, 0, 0.1, 3, dict(obs_type="full_state")),
    ("allegro_hand", 40, 1, 1.0, 2, dict(obs_type="full", asymmetric_obs=True)),
    ("allegro_hand", 40, 2, 0.1, 2, dict(obs_type="full_no_vel", max_consecutive_successes=3, goal_reset_prob=0.2)),
    ("franka_reach", 64, 0, 0.1, 3, {}),
    ("franka_reach", 32, 1, 1.0, 2, {}),                          # first step resets every env (vec_task.py:317)
    ("franka_reach", 32, 2, 0.3, 2, dict(ragged_flags=True)),     # AND filter: partially flagged envs do not reset
    ("franka_reach", 24, 3, 0.2, 2, dict(num_agents=3)),          # numAgents override (cfg/task/FrankaReachMA.yaml:9)
    ("franka_collect", 64, 0, 0.1, 3, {}),
    ("franka_collect", 32, 1, 1.0, 2, {}),
    ("franka_combine", 64, 0, 0.1, 3, {}),
    ("franka_combine", 32, 1, 1.0, 2, {}),
    ("franka_pp", 64, 0, 0.1, 3, {}),
    ("franka_pp", 32, 1, 1.0, 2, {}),
    ("franka_pp", 32, 2, 0.3, 2, dict(ragged_flags=True)),
]


def generate(task, n, seed, reset_prob, steps, **kw):
    st = syn.make_state(task, n, seed=seed, reset_prob=reset_prob, **kw)
    cfg = CFGS[task]
    if task.startswith("franka_"):
        cfg = syn.franka_ma_cfg(task[len("franka_"):], int(st["cfg_num_agents"]), int(st["cfg_num_targets"]))
    obj = ref_harness.BUILDERS[task](st, cfg)
    rng = np.random.Generator(np.random.PCG64(seed + 7919))
    blob = {"meta_task": task, "meta_num_envs": n, "meta_seed": seed, "meta_steps": steps,
            "meta_reset_prob": reset_prob}
    for k, v in st.items():
        blob["init/" + k] = v
    for s in range(steps):
        if s > 0:
            # sim-owned tensors advance; path-owned buffers carry over from the reference
            for name in ref_harness.OUTPUTS[task]:
                st[name] = getattr(obj, name).detach().numpy().copy()
            syn.advance_physics(task, st, rng)
        for name in ref_harness.SIM_INPUTS[task]:
            blob[f"step{s}/in/{name}"] = st[name].copy()
        out, rands, calls = ref_harness.run_reference_step(task, obj, st, cfg, torch_seed=1000 + s)
        for name, v in out.items():
            blob[f"step{s}/out/{name}"] = v
        if isinstance(rands, dict):
            for name, r in rands.items():
                blob[f"step{s}/rand{name}"] = r
        else:
            for i, r in enumerate(rands):
                blob[f"step{s}/rand{i}"] = r
        for i, (cname, args) in enumerate(calls):
            blob[f"step{s}/call{i}/name"] = cname
            idx = [a for a in args if isinstance(a, np.ndarray) and a.dtype == np.int32]
            if idx:
                blob[f"step{s}/call{i}/indices"] = idx[0]
    return blob


def main():
    only = sys.argv[1:]
    for fx in FIXTURES:
        task, n, seed, p, steps = fx[:5]
        kw = fx[5] if len(fx) > 5 else {}
        if only and task not in only:
            continue
        blob = generate(task, n, seed, p, steps, **kw)
        path = os.path.join(HERE, f"{task}_n{n}_s{seed}.npz")
        np.savez_compressed(path, **blob)
        print(f"wrote {path}  ({os.path.getsize(path) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main()
