"""Drive Anymal's post-physics step through the CPU oracle (host pointers) or the CUDA library
(device pointers) from the same IgeAnymalParams / IgeAnymalBuffers; fixtures from tests/golden/gen_golden_anymal.py."""
import glob
import os

import numpy as np
import torch

import harness as H

FLOAT_OUT = ["obs_buf", "rew_buf", "dof_state", "commands", "obs_clamped"]


def golden_files():
    return sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "anymal_n*.npz")))


def cfg_of(blob):
    return {k[4:]: blob[k] for k in blob if k.startswith("cfg_")}


def make_params(cfg, n, rand_ptrs=None, seed=0, step_ptr=None, dt=0.02):
    abi = H.pkg._abi
    p = abi.IgeAnymalParams()
    p.num_envs, p.num_bodies, p.base_index = n, int(cfg["num_bodies"]), int(cfg["base_index"])
    knees = [int(k) for k in np.asarray(cfg["knee_indices"]).tolist()]
    p.num_knees = len(knees)
    abi.fill_array(p.knee_indices, knees)
    for k in ("lin_vel_scale", "ang_vel_scale", "dof_pos_scale", "dof_vel_scale", "clip_obs"):
        setattr(p, k, float(cfg[k]))
    p.rew_lin_vel_xy = float(cfg["rew_lin_vel_xy"]) * dt     # anymal.py:99-100 (Python double product, then fp32)
    p.rew_ang_vel_z = float(cfg["rew_ang_vel_z"]) * dt
    p.rew_torque = float(cfg["rew_torque"]) * dt
    p.max_episode_length = float(cfg["max_episode_length"])
    for k in ("command_x_range", "command_y_range", "command_yaw_range", "default_dof_pos"):
        abi.fill_array(getattr(p, k), np.asarray(cfg[k], np.float64).tolist())
    if rand_ptrs is None:
        p.rand_mode, p.seed, p.step = abi.IGE_RAND_PHILOX, seed, step_ptr
    else:
        p.rand_mode = abi.IGE_RAND_INJECTED
        for i, ptr in enumerate(rand_ptrs):
            p.u[i] = ptr
    return p


class State:
    """Buffers of one env instance on `device` ("cpu" -> oracle, "cuda:0" -> library)."""

    def __init__(self, n, num_bodies, device):
        f = dict(device=device, dtype=torch.float32)
        self.n, self.device = n, device
        self.root_states, self.dof_state = torch.zeros((n, 13), **f), torch.zeros((n, 12, 2), **f)
        self.contact_forces, self.torques = torch.zeros((n, num_bodies, 3), **f), torch.zeros((n, 12), **f)
        self.actions, self.commands = torch.zeros((n, 12), **f), torch.zeros((n, 3), **f)
        self.obs_buf, self.rew_buf, self.obs_clamped = torch.zeros((n, 48), **f), torch.zeros((n,), **f), torch.zeros((n, 48), **f)
        self.reset_buf = torch.zeros((n,), device=device, dtype=torch.int64)
        self.progress_buf = torch.zeros((n,), device=device, dtype=torch.int64)
        self.time_outs = torch.zeros((n,), device=device, dtype=torch.uint8)
        self.reset_ids = torch.full((max(n, 1),), -1, device=device, dtype=torch.int32)
        self.reset_count = torch.zeros((1,), device=device, dtype=torch.int32)
        words = 4 + (n + 127) // 128 + 8
        self.scratch = torch.zeros((words,), device=device, dtype=torch.int32)
        self.rng_step = torch.zeros((1,), device=device, dtype=torch.int64)

    def load(self, **arrays):
        for k, v in arrays.items():
            getattr(self, k).copy_(torch.from_numpy(np.ascontiguousarray(v)))

    def buffers(self):
        b = H.pkg._abi.IgeAnymalBuffers()
        for k in ("root_states", "dof_state", "contact_forces", "torques", "actions", "commands", "obs_buf", "rew_buf",
                  "reset_buf", "progress_buf", "obs_clamped", "reset_ids", "reset_count", "scratch"):
            setattr(b, k, getattr(self, k).data_ptr())
        b.timeout_buf = self.time_outs.data_ptr()
        return b

    def step(self, cfg, rands=None, seed=0):
        """rands: list of 5 numpy arrays (injected mode) or None (Philox)."""
        keep = None
        if rands is not None:
            keep = [torch.from_numpy(np.ascontiguousarray(r, np.float32)).to(self.device) for r in rands]
            ptrs = [t.data_ptr() if t.numel() else None for t in keep]
            p = make_params(cfg, self.n, rand_ptrs=ptrs)
        else:
            p = make_params(cfg, self.n, seed=seed, step_ptr=self.rng_step.data_ptr())
        b = self.buffers()
        if self.device == "cpu":
            H.oracle.call("anymal_post_physics", p, b)
        else:
            H.pkg._lib.call("ige_anymal_post_physics", p, b, stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
        del keep
        k = int(self.reset_count)
        return self.reset_ids[:k].cpu().numpy()

    def outputs(self):
        out = {k: getattr(self, k).cpu().numpy() for k in FLOAT_OUT + ["reset_buf", "progress_buf", "time_outs"]}
        return out


def run_fixture(blob, device):
    """Replay a golden fixture; yields (step, reset ids, outputs, reference outputs)."""
    cfg = cfg_of(blob)
    n, steps = int(blob["meta_num_envs"]), int(blob["meta_steps"])
    st = State(n, int(cfg["num_bodies"]), device)
    st.load(reset_buf=blob["init/reset_buf"], progress_buf=blob["init/progress_buf"], commands=blob["init/commands"])
    for s in range(steps):
        ins = {k: blob[f"step{s}/in/{k}"] for k in ("root_states", "dof_state", "contact_forces", "torques")}
        st.load(actions=blob[f"step{s}/out/actions"], **ins)       # self.actions = clamped clone (anymal.py:227)
        rands = [blob.get(f"step{s}/rand{i}", np.zeros((0,), np.float32)) for i in range(5)]
        ids = st.step(cfg, rands)
        ref = {k: blob[f"step{s}/out/{k}"] for k in FLOAT_OUT + ["reset_buf", "progress_buf", "time_outs", "reset_env_ids"]}
        yield s, ids, st.outputs(), ref


def compare(got, ref, what, bitexact=("dof_state", "commands")):
    """flags / counters bit-exact; reset rows (dof_state, commands) bit-exact; obs / rewards within 1e-5 relative."""
    for k in ("reset_buf", "progress_buf"):
        H.assert_equal(f"{what} {k}", got[k], ref[k])
    H.assert_equal(f"{what} time_outs", got["time_outs"].astype(bool), np.asarray(ref["time_outs"]).astype(bool))
    for k in FLOAT_OUT:
        g, r = got[k].astype(np.float64), np.asarray(ref[k], np.float64).reshape(got[k].shape)
        if k in bitexact:
            H.assert_equal(f"{what} {k}", got[k], np.asarray(ref[k], np.float32).reshape(got[k].shape))
        err = np.abs(g - r) / np.maximum(1.0, np.abs(r))
        assert err.max() <= H.RTOL, f"{what} {k}: max rel err {err.max():.3e}"
