"""CPU-only: the C-ABI library builds, loads and exports every symbol include/ige_b200.h declares;
the ctypes mirror matches the C struct layout; bad arguments are rejected with IGE_E* codes and a
CPU tensor never reaches a kernel (no fallback)."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import pytest
import torch

import harness as H

ROOT = H.ROOT
pkg = H.pkg


@pytest.fixture(scope="module")
def lib():
    from isaacgymenvs_ma_b200 import build
    build.build()
    return pkg._lib.lib()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "ige_b200.h")).read()
    declared = set(re.findall(r"^\s*(?:int|void|uint64_t|const char \*)\s*\*?\s*(ige_\w+)\s*\(", hdr, flags=re.M))
    assert {"ige_version", "ige_ant_post_physics", "ige_humanoid_post_physics", "ige_cartpole_post_physics",
            "ige_compact_resets"} <= declared
    for sym in declared:
        assert hasattr(lib, sym), f"{sym} declared in ige_b200.h but not exported"
    assert set(pkg._lib.exported_symbols()) == declared
    assert lib.ige_version() == 1


def test_ctypes_layout_matches_header():
    structs = [n for n in dir(pkg._abi) if n.startswith("Ige") and isinstance(getattr(pkg._abi, n), type)]
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "ige_b200.h"', "int main(void){"]
    for s in structs:
        src.append(f'printf("{s} %zu\\n", sizeof({s}));')
        for fname, _ in getattr(pkg._abi, s)._fields_:
            src.append(f'printf("{s}.{fname} %zu\\n", offsetof({s}, {fname}));')
    src.append("return 0;}")
    with tempfile.TemporaryDirectory() as td:
        c = os.path.join(td, "probe.c")
        open(c, "w").write("\n".join(src))
        exe = os.path.join(td, "probe")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe], check=True)
        out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout
    got = dict(line.split() for line in out.strip().splitlines())
    for s in structs:
        cls = getattr(pkg._abi, s)
        assert int(got[s]) == C.sizeof(cls), s
        for fname, _ in cls._fields_:
            assert int(got[f"{s}.{fname}"]) == getattr(cls, fname).offset, f"{s}.{fname}"


def test_bad_arguments_rejected(lib):
    p, b = pkg._abi.IgeLocomotionParams(), pkg._abi.IgeLocomotionBuffers()
    p.num_envs, p.num_dof, p.num_sensor_floats = 4, 8, 24
    with pytest.raises(pkg._lib.IgeError, match="NULL"):
        pkg._lib.call("ige_ant_post_physics", p, b)
    p.num_dof = 9
    with pytest.raises(pkg._lib.IgeError, match="num_dof"):
        pkg._lib.call("ige_ant_post_physics", p, b)
    cp, cb = pkg._abi.IgeCartpoleParams(), pkg._abi.IgeCartpoleBuffers()
    cp.num_envs = 4
    cb.dof_state = 8  # misaligned, non-NULL
    with pytest.raises(pkg._lib.IgeError, match="aligned"):
        pkg._lib.call("ige_cartpole_post_physics", cp, cb)


def test_no_cpu_fallback():
    st = pkg.synthetic.make_state("ant", 64, seed=0)
    env = H.build("ant", st, "cpu")
    with pytest.raises(pkg._lib.IgeError, match="no CPU path"):
        env.post_physics_step()


def test_buffers_match_vec_task_contract():
    # vec_task.py:302-326
    env = pkg.make(0, "Ant", 32, "cpu")
    assert env.obs_buf.shape == (32, 60) and env.obs_buf.dtype == torch.float32
    assert env.rew_buf.shape == (32,) and env.rew_buf.dtype == torch.float32
    for name in ("reset_buf", "progress_buf", "randomize_buf", "timeout_buf"):
        t = getattr(env, name)
        assert t.shape == (32,) and t.dtype == torch.long, name
    hum = pkg.make(0, "Humanoid", 16, "cpu")
    assert hum.obs_buf.shape == (16, 108)
    cp = pkg.make(0, "Cartpole", 16, "cpu")
    assert cp.obs_buf.shape == (16, 4) and cp.dof_state.shape == (32, 2)
    assert cp.dof_pos.data_ptr() == cp.dof_state.data_ptr()     # strided views, cartpole.py:53-54


def test_bad_arguments_rejected_8f_entry_points(lib):
    """The §8f entry points validate before launching (no GPU needed): negative IGE_E* codes -> IgeError."""
    a = pkg._abi.IgeRolloutStepArgs()
    a.num_rows, a.num_agents, a.meter_max_size = 8, 2, 100
    with pytest.raises(pkg._lib.IgeError, match="non-NULL"):
        pkg._lib.call("ige_ma_rollout_step", a)
    a.num_agents = 0
    with pytest.raises(pkg._lib.IgeError, match="num_agents"):
        pkg._lib.call("ige_ma_rollout_step", a)
    o = pkg._abi.IgeFrankaOscArgs()
    o.num_envs, o.num_agents, o.num_actions, o.num_bodies = 4, 2, 5, 24
    with pytest.raises(pkg._lib.IgeError, match="num_actions"):
        pkg._lib.call("ige_franka_ma_pre_physics", o)
    o.num_actions, o.eef_body_start, o.eef_body_step = 7, 20, 11
    with pytest.raises(pkg._lib.IgeError, match="out of range"):
        pkg._lib.call("ige_franka_ma_pre_physics", o)
    p, b = pkg._abi.IgeAnymalParams(), pkg._abi.IgeAnymalBuffers()
    p.num_envs, p.num_bodies, p.num_knees = 4, 17, 9
    with pytest.raises(pkg._lib.IgeError, match="num_knees"):
        pkg._lib.call("ige_anymal_post_physics", p, b)
    p.num_knees = 4
    with pytest.raises(pkg._lib.IgeError, match="NULL"):
        pkg._lib.call("ige_anymal_post_physics", p, b)
    assert lib.ige_anymal_scratch_words(16384) == 4 + 128 and lib.ige_rollout_scratch_words(16384) == 4 + 16
    with pytest.raises(pkg._lib.IgeError, match="no CPU fallback"):
        pkg.rollout.RolloutBookkeeperMA(16, 2, 4, "cpu")


def test_every_task_of_the_map_constructs_with_the_reference_shapes():
    """isaacgym_task_map names on the path (tasks/__init__.py:67-100) -> mirrors whose contract buffers have the
    reference's shapes (numObservations of each task's __init__, agent-axis rows = env * A + agent)."""
    want = {"Cartpole": 4, "Ant": 60, "Humanoid": 108, "ShadowHand": 211, "AllegroHand": 88, "Anymal": 48,
            "BallBalance": 24, "Quadcopter": 21, "Ingenuity": 13, "FrankaReach": 14, "FrankaCubeStack2": 21,
            "FrankaCubeStack": 19, "FrankaCabinet": 23, "FrankaReachMA": 19, "FrankaCollectMA": 28,
            "FrankaCombineMA": 48, "FrankaPPMA": 50}
    assert set(want) == set(pkg.tasks.task_map)
    for task, nobs in want.items():
        env = pkg.make(0, task, 32, "cpu", graphics_device_id=0, headless=True, multi_gpu=False,
                       virtual_screen_capture=False, force_render=False)      # the reference's own keyword set
        rows = 32 * env.num_agents
        assert env.obs_buf.shape == (rows, nobs), task
        assert env.reset_buf.shape == (rows,) and env.reset_buf.dtype == torch.long, task
        assert env.get_number_of_agents() == (2 if task.endswith("MA") else 1), task
        obs, ids = env.reset_done()
        assert obs["obs"].shape == (rows, nobs)


def test_oracle_philox_is_a_pure_function_of_seed_env_step():
    """Production reset noise (Philox4x32-10 keyed by seed, counter = env, draw, step): two oracle runs from the same
    state agree bit for bit, the step word advances once per call, a different seed gives different rows."""
    import franka_sa_harness as F
    st = pkg.synthetic.make_state("franka_sa_stack", 257, seed=3, reset_prob=0.5)
    outs = []
    for seed in (5, 5, 6):
        s = F.State("franka_sa_stack", 257, "cpu")
        s.load_state(st)
        s.post(None, seed=seed)
        assert int(s.rng_step) == 1
        outs.append(s.get(("root_state", "dof_state", "init_cubeA_state")))
    for k in outs[0]:
        assert (outs[0][k] == outs[1][k]).all()
    assert not (outs[0]["dof_state"] == outs[2]["dof_state"]).all()
    # FrankaCubeStack: the re-sampled cube A never ends up closer to cube B than the collision-free distance (:550-585)
    import numpy as np
    g = s.get(("init_cubeA_state", "init_cubeB_state"))
    flagged = st["reset_buf"] != 0
    d = np.linalg.norm(g["init_cubeA_state"][flagged, :2] - g["init_cubeB_state"][flagged, :2], axis=1)
    assert flagged.sum() > 50 and (d >= np.float32(0.12) * np.float32(np.sqrt(2)) - 1e-6).all()
