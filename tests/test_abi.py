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
