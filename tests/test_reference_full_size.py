"""Full-size pinning of the oracle ("goldens are small; full-size configs are CUDA-vs-oracle
only").  The committed goldens are N <= 512; here the UNMODIFIED reference classes are stepped at BASELINE.json's own sizes
(Ant 16384, Humanoid 32768, Franka PP / Reach 8192 x 2, ShadowHand 16384) and compared with the C oracle on the same state,
live -- no fixture files.  Together with the -m gpu tests (CUDA == oracle at the same sizes) this closes the chain
reference == oracle == CUDA at full size.

Needs the reference sources: /root/reference in the build container, or the copy tools/vendor_reference.py stages under
oracle/_ref/ (which travels to the GPU box).  CPU only."""
import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.needs_reference


def _cfg(task, st):
    syn = H.pkg.synthetic
    if task.startswith("franka_"):
        return syn.franka_ma_cfg(task[len("franka_"):], int(st["cfg_num_agents"]), int(st["cfg_num_targets"]))
    return {"ant": syn.ANT_CFG, "humanoid": syn.HUMANOID_CFG, "shadow_hand": syn.SHADOW_HAND_CFG,
            "allegro_hand": syn.ALLEGRO_HAND_CFG, "cartpole": syn.CARTPOLE_CFG}[task]


@pytest.mark.parametrize("task,n", [("cartpole", 512), ("ant", 16384), ("humanoid", 32768), ("franka_pp", 8192),
                                    ("franka_reach", 8192), ("shadow_hand", 16384), ("allegro_hand", 16384)])
def test_oracle_matches_reference_at_baseline_size(task, n):
    import ref_harness
    st = H.pkg.synthetic.make_state(task, n, seed=17, reset_prob=0.02)
    cfg = _cfg(task, st)
    ref_obj = ref_harness.BUILDERS[task]({k: v.copy() for k, v in st.items()}, cfg)
    mirror = H.build(task, {k: v.copy() for k, v in st.items()}, "cpu")
    rng = np.random.Generator(np.random.PCG64(5))
    cur = {k: v.copy() for k, v in st.items()}
    for step in range(2):
        if step > 0:
            for name in ref_harness.OUTPUTS[task]:
                cur[name] = getattr(ref_obj, name).detach().numpy().copy()
            H.pkg.synthetic.advance_physics(task, cur, rng)
        names = [k for k in ref_harness.SIM_INPUTS[task]]
        H.pkg.factory.load_sim_inputs(mirror, H.TASK_NAME[task], {k: cur[k] for k in names if k in cur})
        out, rands, _calls = ref_harness.run_reference_step(task, ref_obj, cur, cfg, torch_seed=300 + step)
        ids = H.oracle_step(task, mirror, rands)
        H.assert_equal(f"{task} step{step} reset_env_ids", ids.numpy(), out["reset_env_ids"])
        H.compare_outputs(task, mirror, out, bitexact=H.BITEXACT_FLOAT[task])
        frac = H.bitexact_fraction(H.get(mirror, "obs_buf"), out["obs_buf"])
        assert frac > 0.97, (task, step, frac)
    assert int((H.get(mirror, "reset_buf") != 0).sum()) > 0


@pytest.mark.parametrize("task", ["franka_sa_reach", "franka_sa_stack2", "franka_sa_stack", "franka_cabinet",
                                  "ball_balance", "quadcopter", "ingenuity"])
def test_oracle_matches_reference_8f_row2_at_16384(task):
    """SURVEY 8f row 2 tasks at 16384 envs: the fixture generators (tests/golden/gen_golden_*.py) step the unmodified
    reference class in memory, the oracle replays the trajectory."""
    import os
    import sys
    sys.path.insert(0, os.path.join(H.ROOT, "tests", "golden"))
    if task.startswith("franka"):
        import franka_sa_harness as F
        import gen_golden_franka_sa as G
        blob, resets = G.generate(task, 16384, 23, 0.02, 2)
        assert sum(resets) > 100
        for s, got, ref in F.run_fixture(blob, task, "cpu"):
            F.compare(got, ref, f"{task} 16384 step{s}")
            assert H.bitexact_fraction(got["obs_buf"], ref["obs_buf"]) > 0.97
    else:
        import gen_golden_small as G
        import small_harness as S
        from test_oracle_golden import SMALL_EXACT
        blob = G.generate(task, 16384, 23, 0.02, 2)
        for s, phase, got, ref in S.run_fixture(blob, task, "cpu"):
            S.compare(got, ref, f"{task} 16384 step{s} {phase}", exact_float=SMALL_EXACT[task])


def test_oracle_matches_reference_anymal_osc_rollout_at_size():
    """The remaining 8f rows at C4-sized inputs: Anymal 16384 envs (VecTask.step), the OSC pre_physics_step of FrankaPPMA at
    8192 x 2 agent rows, A2CAgentMA.play_steps on 8192 x 2 rows -- reference run in memory, oracle replays."""
    import os
    import sys
    sys.path.insert(0, os.path.join(H.ROOT, "tests", "golden"))
    import anymal_harness as A
    import gen_golden_anymal as GA
    blob = GA.generate(16384, 31, 0.02, 2)
    for s, ids, got, ref in A.run_fixture(blob, "cpu"):
        H.assert_equal("anymal reset_env_ids", ids, ref["reset_env_ids"])
        A.compare(got, ref, f"anymal 16384 step{s}")
    import gen_golden_osc as GO
    import osc_harness as O
    blob = GO.generate("pp", 8192, 2, 7)
    got, truth = O.oracle_osc(blob), O.float64_truth(blob)
    H.assert_equal("osc actions", got["out_actions"], blob["out_actions"])
    H.assert_equal("osc pos_control", got["out_pos_control"], blob["out_pos_control"])
    assert O.torque_error(got, truth) <= 2.0 * O.torque_error(blob, truth) + 1e-5    # as close to float64 as the reference's fp32
    over, rows = O.rows_over(got, blob)
    print(f"OSC pp 8192x2: {over} of {rows} agent rows beyond 1e-5 of the reference's fp32 torques")
    import gen_golden_rollout as GR
    import rollout_harness as R
    cls = GR.load_reference_agent()
    case = dict(GR.CASES[0])
    case.pop("name")
    case.update(num_envs=8192, A=2)
    blob = GR.run_case(cls, **case)
    R.compare(R.oracle_rollout(blob), blob, "rollout 8192x2")
