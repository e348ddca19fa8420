"""Pin the CPU oracle (oracle/ige_oracle.c) against fixtures minted from the unmodified reference
(tests/golden/gen_golden.py).  Runs anywhere (no GPU, no reference tree)."""
import os

import numpy as np
import pytest

import harness as H


def _run_fixture(task, path):
    blob = H.load_golden(path)
    st = H.init_state(blob)
    obj = H.build(task, st, "cpu")
    steps = int(blob["meta_steps"])
    worst = {}
    for s in range(steps):
        H.pkg.factory.load_sim_inputs(obj, H.TASK_NAME[task], H.step_inputs(blob, s))
        ids = H.oracle_step(task, obj, H.step_rands(blob, s))
        ref = H.step_outputs(blob, s)
        H.assert_equal("reset_env_ids", ids.numpy(), ref["reset_env_ids"])
        if task not in H.HAND_TASKS and task not in H.MA_TASKS:   # (actor-id lists: checked in the GPU tests)
            for cname, idx in H.step_call_indices(blob, s):
                H.assert_equal(f"{cname} indices", ids.numpy(), idx)
        H.compare_outputs(task, obj, ref, bitexact=H.BITEXACT_FLOAT[task])
        for name in H.FLOAT_OUT[task]:
            worst[name] = min(worst.get(name, 1.0), H.bitexact_fraction(H.get(obj, name), ref[name]))
    return worst


@pytest.mark.parametrize("task", ["cartpole", "ant", "humanoid", "shadow_hand", "allegro_hand"] + H.MA_TASKS)
def test_oracle_matches_reference_golden(task):
    files = H.golden_files(task)
    assert files, f"no golden fixtures for {task}"
    for path in files:
        worst = _run_fixture(task, path)
        # the restatement is op-for-op: everything but libm-vs-sleef transcendentals is bit-exact
        # (ShadowHand's reward goes through asin(): libm and torch's vectorised asin differ in the last ulp)
        assert worst["rew_buf"] > (0.5 if task in H.HAND_TASKS else 0.9), worst


def test_oracle_rollout_matches_reference_play_steps():
    """the oracle's restatement of A2CAgent_MA.py:35-67 against what the reference's own
    play_steps wrote on a scripted environment (tests/golden/gen_golden_rollout.py)."""
    import rollout_harness as R
    files = R.golden_files()
    assert len(files) >= 4
    for path in files:
        blob = H.load_golden(path)
        R.compare(R.oracle_rollout(blob), blob, os.path.basename(path))


def test_oracle_osc_matches_reference_pre_physics():
    """the oracle's restatement of Franka*MA.pre_physics_step + _compute_osc_torques against the
    reference's own output (tests/golden/gen_golden_osc.py).  Clamped actions and gripper targets bit-exact.  The
    torques go through two matrix inverses (torch.inverse = LAPACK LU, here Gauss-Jordan): fp32 rounding is amplified
    by the condition number, so the bar is 1e-4 * max(1, |ref|), AND the oracle must be as close to a float64
    evaluation of the same formula as the reference's own fp32 result is (within 2x + 1e-5)."""
    import osc_harness as O
    files = O.golden_files()
    assert len(files) >= 4
    for path in files:
        blob = H.load_golden(path)
        got, truth = O.oracle_osc(blob), O.float64_truth(blob)
        H.assert_equal("actions", got["out_actions"], blob["out_actions"])
        H.assert_equal("pos_control", got["out_pos_control"], blob["out_pos_control"])
        assert O.torque_error(got, blob) <= 1e-4, (path, O.torque_error(got, blob))
        over, rows = O.rows_over(got, blob)
        print(f"{os.path.basename(path)}: {over} of {rows} agent rows beyond 1e-5 of the reference's fp32 torques")
        assert O.torque_error(got, truth) <= 2.0 * O.torque_error(blob, truth) + 1e-5
        # untouched columns (6 of each arm, 7..8) keep their previous value: franka_PPMA.py:944 assigns [:, :6]
        eff, ref = got["out_effort_control"].reshape(-1, 9), blob["out_effort_control"].reshape(-1, 9)
        H.assert_equal("untouched effort columns", eff[:, 6:], ref[:, 6:])


def test_oracle_anymal_matches_reference_golden():
    """the oracle against the reference's own VecTask.step outputs
    (tests/golden/gen_golden_anymal.py): flags / progress / reset ids / reset rows bit-exact, obs and rewards 1e-5."""
    import anymal_harness as A
    files = A.golden_files()
    assert len(files) >= 3
    for path in files:
        blob = H.load_golden(path)
        for s, ids, got, ref in A.run_fixture(blob, "cpu"):
            H.assert_equal("reset_env_ids", ids, ref["reset_env_ids"])
            A.compare(got, ref, f"{os.path.basename(path)} step{s}")
            assert H.bitexact_fraction(got["obs_buf"], ref["obs_buf"]) > 0.98


# reset rows, targets, thrusts and forces are plain copies / a few fp32 ops of injected draws: bit-exact -- except
# BallBalance's ball rows, which go through cos / sin of the drawn direction (torch CPU = Sleef, libm and CUDA differ
# from it in the last ulp)
SMALL_EXACT = {t: tuple(n for n in ("root_states", "root_tensor", "dof_states", "dof_position_targets", "thrusts", "forces",
                                    "target_root_positions") if not (t == "ball_balance" and n == "root_states"))
               for t in ("ball_balance", "quadcopter", "ingenuity")}


@pytest.mark.parametrize("task", ["ball_balance", "quadcopter", "ingenuity"])
def test_oracle_small_tasks_match_reference_golden(task):
    """SURVEY.md 8f row 2 (BallBalance / Quadcopter / Ingenuity): the oracle's pre_physics and post_physics halves
    against the reference's own VecTask.step (tests/golden/gen_golden_small.py, gym.simulate hooked to snapshot the
    pre-physics writes): ids / flags / counters and every reset row bit-exact, obs and rewards within 1e-5."""
    import small_harness as S
    files = S.golden_files(task)
    assert len(files) >= 3
    resets = timeouts = 0
    for path in files:
        blob = H.load_golden(path)
        for s, phase, got, ref in S.run_fixture(blob, task, "cpu"):
            S.compare(got, ref, f"{os.path.basename(path)} step{s} {phase}", exact_float=SMALL_EXACT[task])
            if phase == "pre":
                resets += len(got["reset_env_ids"])
            else:
                timeouts += int(got["time_outs"].sum())
                assert H.bitexact_fraction(got["obs_buf"], ref["obs_buf"]) > 0.98
    assert resets > 50 and timeouts > 0


@pytest.mark.parametrize("task", ["franka_sa_reach", "franka_sa_stack2", "franka_sa_stack", "franka_cabinet"])
def test_oracle_franka_sa_matches_reference_golden(task):
    """SURVEY.md 8f row 2 (FrankaReach / FrankaCubeStack2 / FrankaCubeStack / FrankaCabinet): the oracle against the
    reference's own VecTask.step (tests/golden/gen_golden_franka_sa.py): ids / flags / counters, the reset DOF rows,
    control targets and prop rows bit-exact; cube rows (sinf / cosf of the drawn yaw), observations, rewards and the
    reward-component means within 1e-5."""
    import franka_sa_harness as F
    files = F.golden_files(task)
    assert len(files) >= 3
    resets = timeouts = dones = 0
    for path in files:
        blob = H.load_golden(path)
        for s, got, ref in F.run_fixture(blob, task, "cpu"):
            F.compare(got, ref, f"{os.path.basename(path)} step{s}")
            resets += len(got["reset_env_ids"])
            timeouts += int(got["time_outs"].sum())
            dones += int((got["reset_buf"] != 0).sum())
            assert H.bitexact_fraction(got["obs_buf"], ref["obs_buf"]) > 0.98
            assert H.bitexact_fraction(got["rew_buf"], ref["rew_buf"]) > (0.7 if task == "franka_sa_stack" else 0.9)   # tanhf: libm vs Sleef
    assert resets > 40 and timeouts > 0 and dones >= timeouts
