"""Drive BallBalance / Quadcopter / Ingenuity through the CPU oracle (host pointers) or the CUDA
library (device pointers) from the same Args struct; fixtures from tests/golden/gen_golden_small.py.

These tasks reset in PRE-physics position, so a control step is: pre_physics (reset_idx + action mapping) ->
[simulator: new root / DOF / sensor values are loaded] -> post_physics (obs, reward, done)."""
import glob
import os

import numpy as np
import torch

import harness as H

small = H.pkg.tasks.small
syn = H.pkg.synthetic
CFG = {"ball_balance": syn.BALL_BALANCE_CFG, "quadcopter": syn.QUADCOPTER_CFG, "ingenuity": syn.INGENUITY_CFG}
TASKS = tuple(CFG)
# tensors rewritten by pre_physics (compared after the pre call) / by post_physics
PRE_OUT = {
    "ball_balance": ("root_states", "dof_states", "dof_position_targets"),
    "quadcopter": ("root_states", "dof_states", "dof_position_targets", "thrusts", "forces"),
    "ingenuity": ("root_tensor", "dof_states", "target_root_positions", "thrusts", "forces"),
}
POST_OUT = ("obs_buf", "rew_buf", "obs_clamped")
# what the simulator rewrites between the two calls
PHYS = {"ball_balance": ("root_states", "dof_states", "sensor_tensor"), "quadcopter": ("root_states", "dof_states"),
        "ingenuity": ("root_tensor", "dof_states")}


def golden_files(task):
    return sorted(glob.glob(os.path.join(H.GOLDEN_DIR, f"{task}_n*.npz")))


def exclusive_rank(flags):
    f = (np.asarray(flags) != 0).astype(np.int64)
    return (np.cumsum(f) - f).astype(np.int32)


class State:
    """All buffers of one task instance on `device` ("cpu" -> oracle, "cuda:0" -> library)."""

    def __init__(self, task, n, device, cfg=None):
        self.task, self.n, self.device, self.cfg = task, n, device, dict(cfg or CFG[task])
        f = dict(device=device, dtype=torch.float32)
        for name, shape in small.TENSORS[task].items():
            setattr(self, name, torch.zeros((n,) + shape, **f))
        nobs = small.NUM_OBS[task]
        self.obs_buf, self.obs_clamped, self.rew_buf = torch.zeros((n, nobs), **f), torch.zeros((n, nobs), **f), torch.zeros(n, **f)
        self.reset_buf = torch.zeros(n, device=device, dtype=torch.int64)
        self.progress_buf = torch.zeros(n, device=device, dtype=torch.int64)
        self.time_outs = torch.zeros(n, device=device, dtype=torch.uint8)
        self.reset_ids = torch.full((max(n, 1),), -1, device=device, dtype=torch.int32)
        self.reset_count = torch.zeros(1, device=device, dtype=torch.int32)
        self.rank = torch.zeros(max(n, 1), device=device, dtype=torch.int32)
        self.target_rank = torch.zeros(max(n, 1), device=device, dtype=torch.int32)
        self.actor_dirty = torch.zeros(max(2 * n, 1), device=device, dtype=torch.int64)
        self.scratch = torch.zeros(H.pkg._abi.reset_list_scratch_words(max(n, 1)), device=device, dtype=torch.int32)
        self.rng_step = torch.zeros(1, device=device, dtype=torch.int64)
        self.emit_list = True

    def load(self, **arrays):
        for k, v in arrays.items():
            t = getattr(self, k)
            t.copy_(torch.from_numpy(np.ascontiguousarray(v)).view_as(t))

    def load_state(self, st):
        self.load(**{k: v for k, v in st.items() if hasattr(self, k)})
        c = float(self.cfg.get("clip_actions", 1.0))
        self.actions.clamp_(-c, c)                       # vec_task.py:373 (pre_physics_step receives clamped actions)

    def args(self, rands=None, seed=0):
        a = small.fill_scalars(self.task, small.ARGS[self.task](), self.cfg, self.n)
        for name in small.TENSORS[self.task]:
            setattr(a, name, getattr(self, name).data_ptr())
        for name in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "obs_clamped"):
            setattr(a, name, getattr(self, name).data_ptr())
        a.timeout_buf = self.time_outs.data_ptr()
        if self.emit_list:
            a.list.ids, a.list.count, a.list.scratch = self.reset_ids.data_ptr(), self.reset_count.data_ptr(), self.scratch.data_ptr()
        if self.task == "ingenuity":
            a.actor_dirty, a.target_rank = self.actor_dirty.data_ptr(), self.target_rank.data_ptr()
        keep = None
        if rands is None:
            a.rand.mode, a.rand.seed, a.rand.step = H.pkg._abi.IGE_RAND_PHILOX, seed, self.rng_step.data_ptr()
        else:
            a.rand.mode, a.rand.rank = H.pkg._abi.IGE_RAND_INJECTED, self.rank.data_ptr()
            keep = [torch.from_numpy(np.ascontiguousarray(r, np.float32)).to(self.device) for r in rands]
            for i, t in enumerate(keep):
                a.rand.u[i] = t.data_ptr() if t.numel() else None
        return a, keep

    def _call(self, which, a):
        if self.device == "cpu":
            H.oracle.call(f"{self.task}_{which}_physics", a)
        else:
            H.pkg._lib.call(f"ige_{self.task}_{which}_physics", a, stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()

    def pre(self, rands=None, seed=0):
        """rands: the task's injected draws (small.DRAW_COLS order) or None (Philox).  Returns the reset ids."""
        if rands is not None:                            # rows of the draws: exclusive ranks, as compact_resets gives
            self.rank[:self.n].copy_(torch.from_numpy(exclusive_rank(self.reset_buf.cpu().numpy())))
            if self.task == "ingenuity":
                self.target_rank[:self.n].copy_(torch.from_numpy(exclusive_rank(self.progress_buf.cpu().numpy() % 500 == 0)))
        a, keep = self.args(rands, seed)
        self._call("pre", a)
        del keep
        return self.reset_ids[:int(self.reset_count)].cpu().numpy()

    def post(self):
        a, _ = self.args(None)
        self._call("post", a)

    def get(self, names):
        return {k: getattr(self, k).cpu().numpy().copy() for k in names}


def draw_shapes(task, k, kt=0):
    return [((kt if (task == "ingenuity" and i < 2) else k), c) for i, c in enumerate(small.DRAW_COLS[task])]


def np_draws(task, rng, k, kt=0):
    return [rng.random(s, dtype=np.float32) for s in draw_shapes(task, k, kt)]


def run_fixture(blob, task, device):
    """Replay a golden fixture; yields (step, phase, got, ref) for phase in ("pre", "post")."""
    n, steps = int(blob["meta_num_envs"]), int(blob["meta_steps"])
    st = State(task, n, device)
    st.load_state({k[5:]: blob[k] for k in blob if k.startswith("init/")})
    for s in range(steps):
        st.load(actions=blob[f"step{s}/actions_clamped"])
        rands = [blob[f"step{s}/rand{i}"] for i in range(len(small.DRAW_COLS[task]))]
        ids = st.pre(rands)
        got = st.get(PRE_OUT[task] + ("reset_buf", "progress_buf"))
        got["reset_env_ids"] = ids
        if task == "ingenuity":
            got["actor_ids"] = np.flatnonzero(st.actor_dirty.cpu().numpy()[:2 * n]).astype(np.int32)
        ref = {k: blob[f"step{s}/mid/{k}"] for k in got}
        yield s, "pre", got, ref
        st.load(**{k: blob[f"step{s}/phys/{k}"] for k in PHYS[task]})
        st.post()
        got = st.get(POST_OUT + ("reset_buf", "progress_buf", "time_outs"))
        ref = {k: blob[f"step{s}/out/{k}"] for k in got}
        yield s, "post", got, ref


def compare(got, ref, what, exact_float=()):
    """ids / flags / counters bit-exact; floats within 1e-5 relative (names in `exact_float`: bit-exact)."""
    for k, g in got.items():
        r = np.asarray(ref[k])
        if g.dtype.kind in "iub" or r.dtype.kind in "iub":
            H.assert_equal(f"{what} {k}", g.astype(np.int64).ravel(), r.astype(np.int64).ravel())
        elif k in exact_float:
            H.assert_equal(f"{what} {k}", g.ravel().view(np.uint32), r.astype(np.float32).ravel().view(np.uint32))
        else:
            H.assert_float_close(f"{what} {k}", g.ravel(), r.astype(np.float32).ravel())
