"""Drive FrankaReach / FrankaCubeStack2 / FrankaCubeStack / FrankaCabinet through the CPU oracle
(host pointers) or the CUDA library (device pointers) from the same Args struct; fixtures from
tests/golden/gen_golden_franka_sa.py."""
import glob
import os

import numpy as np
import torch

import harness as H

fsa = H.pkg.tasks.franka_sa
syn = H.pkg.synthetic
abi = H.pkg._abi
TASKS = ("franka_sa_reach", "franka_sa_stack2", "franka_sa_stack", "franka_cabinet")
PHYS = {t: ("rigid_body_state", "root_state", "dof_state") for t in TASKS[:3]}
PHYS["franka_cabinet"] = ("rigid_body_states", "root_state_tensor", "dof_state")
# tensors post_physics_step rewrites (reset rows) or produces
OUT = {t: ("root_state", "dof_state", "pos_control", "effort_control", "init_cubeA_state", "init_cubeB_state",
           "obs_buf", "rew_buf", "obs_clamped") for t in TASKS[:3]}
OUT["franka_cabinet"] = ("root_state_tensor", "dof_state", "franka_dof_targets", "obs_buf", "rew_buf", "obs_clamped",
                         "franka_grasp_pos", "franka_grasp_rot", "drawer_grasp_pos", "drawer_grasp_rot")
# bit-exact float outputs: everything the reset writes (the draws are injected) except the yaw quaternion (sinf / cosf)
EXACT = ("dof_state", "pos_control", "effort_control", "franka_dof_targets", "root_state_tensor")


def cfg_of(task):
    return dict(syn.FRANKA_CABINET_CFG) if task == "franka_cabinet" else syn.franka_sa_cfg(task[len("franka_sa_"):])


def draw_cols(task):
    return fsa.FrankaCabinet.DRAWS if task == "franka_cabinet" else fsa.DRAW_COLS[task[len("franka_sa_"):]]


def golden_files(task):
    return sorted(glob.glob(os.path.join(H.GOLDEN_DIR, f"{task}_n*.npz")))


def exclusive_rank(flags):
    f = (np.asarray(flags) != 0).astype(np.int64)
    return (np.cumsum(f) - f).astype(np.int32)


class State:
    def __init__(self, task, n, device):
        self.task, self.n, self.device, self.cfg = task, n, device, cfg_of(task)
        c, f = self.cfg, dict(device=device, dtype=torch.float32)
        z = lambda *s: torch.zeros((n,) + s, **f)
        if task == "franka_cabinet":
            self.rigid_body_states, self.root_state_tensor = z(c["num_bodies"], 13), z(c["num_actors"], 13)
            self.default_prop_states, self.dof_state = z(c["num_props"], 13), z(c["num_dofs"], 2)
            self.actions, self.franka_dof_targets = z(9), z(c["num_dofs"])
            self.franka_grasp_pos, self.franka_grasp_rot, self.drawer_grasp_pos, self.drawer_grasp_rot = z(3), z(4), z(3), z(4)
        else:
            self.rigid_body_state, self.root_state, self.dof_state = z(c["num_bodies"], 13), z(c["num_actors"], 13), z(9, 2)
            self.actions, self.pos_control, self.effort_control = z(7), z(9), z(9)
            self.init_cubeA_state, self.init_cubeB_state = z(13), z(13)
            self.reward_components = torch.zeros(abi.IGE_FRANKA_SA_COMPONENTS, **f)
            self.reduce_scratch = torch.zeros(abi.IGE_FRANKA_SA_COMPONENTS + 2, **f)
        nobs = c["num_obs"]
        self.obs_buf, self.obs_clamped, self.rew_buf = z(nobs), z(nobs), z()
        self.reset_buf = torch.zeros(n, device=device, dtype=torch.int64)
        self.progress_buf = torch.zeros(n, device=device, dtype=torch.int64)
        self.time_outs = torch.zeros(n, device=device, dtype=torch.uint8)
        self.reset_ids = torch.full((max(n, 1),), -1, device=device, dtype=torch.int32)
        self.reset_count = torch.zeros(1, device=device, dtype=torch.int32)
        self.rank = torch.zeros(max(n, 1), device=device, dtype=torch.int32)
        self.scratch = torch.zeros(abi.reset_list_scratch_words(max(n, 1)), device=device, dtype=torch.int32)
        self.rng_step = torch.zeros(1, device=device, dtype=torch.int64)

    def load(self, **arrays):
        for k, v in arrays.items():
            t = getattr(self, k)
            t.copy_(torch.from_numpy(np.ascontiguousarray(v)).view_as(t))

    def load_state(self, st):
        self.load(**{k: v for k, v in st.items() if hasattr(self, k) and not k.startswith("cfg_")})
        self.actions.clamp_(-1.0, 1.0)                   # vec_task.py:373

    def args(self, rands=None, seed=0):
        if self.task == "franka_cabinet":
            a = fsa.fill_cabinet_scalars(abi.IgeFrankaCabinetArgs(), self.cfg, self.n)
            names = ("rigid_body_states", "root_state_tensor", "default_prop_states", "dof_state", "actions",
                     "franka_dof_targets", "franka_grasp_pos", "franka_grasp_rot", "drawer_grasp_pos", "drawer_grasp_rot")
        else:
            a = fsa.fill_sa_scalars(abi.IgeFrankaSAArgs(), self.cfg, self.n)
            names = ("rigid_body_state", "root_state", "dof_state", "actions", "pos_control", "effort_control",
                     "init_cubeA_state", "init_cubeB_state", "reward_components", "reduce_scratch")
        for name in names + ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "obs_clamped"):
            setattr(a, name, getattr(self, name).data_ptr())
        a.timeout_buf = self.time_outs.data_ptr()
        a.list.ids, a.list.count, a.list.scratch = self.reset_ids.data_ptr(), self.reset_count.data_ptr(), self.scratch.data_ptr()
        keep = None
        if rands is None:
            a.rand.mode, a.rand.seed, a.rand.step = abi.IGE_RAND_PHILOX, seed, self.rng_step.data_ptr()
        else:
            a.rand.mode, a.rand.rank = abi.IGE_RAND_INJECTED, self.rank.data_ptr()
            keep = [torch.from_numpy(np.ascontiguousarray(r, np.float32)).to(self.device) for r in rands]
            for i, t in enumerate(keep):
                a.rand.u[i] = t.data_ptr() if t.numel() else None
        return a, keep

    def _call(self, entry, a):
        if self.device == "cpu":
            H.oracle.call(entry, a)
        else:
            H.pkg._lib.call("ige_" + entry, a, stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()

    def pre(self):
        if self.task == "franka_cabinet":
            a, _ = self.args(None)
            self._call("franka_cabinet_pre_physics", a)

    def post(self, rands=None, seed=0):
        """rands: injected draws (draw_cols order) or None (Philox).  Returns the reset env ids."""
        if rands is not None:
            self.rank[:self.n].copy_(torch.from_numpy(exclusive_rank(self.reset_buf.cpu().numpy())))
        a, keep = self.args(rands, seed)
        self._call("franka_cabinet_post_physics" if self.task == "franka_cabinet" else "franka_sa_post_physics", a)
        del keep
        return self.reset_ids[:int(self.reset_count)].cpu().numpy()

    def get(self, names):
        return {k: getattr(self, k).cpu().numpy().copy() for k in names}


def np_draws(task, rng, k):
    return [rng.random((k, c), dtype=np.float32) for c in draw_cols(task)]


def run_fixture(blob, task, device):
    """Replay a golden fixture; yields (step, got, ref)."""
    n, steps = int(blob["meta_num_envs"]), int(blob["meta_steps"])
    st = State(task, n, device)
    st.load_state({k[5:]: blob[k] for k in blob if k.startswith("init/")})
    for s in range(steps):
        st.load(actions=blob[f"step{s}/actions_clamped"])
        st.load(**{k: blob[f"step{s}/phys/{k}"] for k in PHYS[task]})
        st.pre()
        if task == "franka_cabinet":
            H.assert_equal(f"{task} step {s} pre franka_dof_targets", st.franka_dof_targets.cpu().numpy().view(np.uint32),
                           blob[f"step{s}/mid/franka_dof_targets"].view(np.uint32))
        rands = [blob[f"step{s}/rand{i}"] for i in range(len(draw_cols(task)))]
        ids = st.post(rands)
        got = st.get(OUT[task] + ("reset_buf", "progress_buf", "time_outs"))
        got["reset_env_ids"] = ids
        if task in ("franka_sa_reach", "franka_sa_stack2"):
            comp = st.reward_components.cpu().numpy()
            names = [str(x) for x in blob[f"step{s}/out/component_names"]]
            idx = fsa.COMPONENT_NAMES[task[len("franka_sa_"):]]
            got["components"] = np.array([comp[idx[nm]] for nm in names], np.float32)
        ref = {k: blob[f"step{s}/out/{k}"] for k in got}
        yield s, got, ref


def compare(got, ref, what, exact_float=EXACT):
    """ids / flags / counters bit-exact; floats within 1e-5 relative (names in `exact_float`: bit-exact)."""
    for k, g in got.items():
        r = np.asarray(ref[k])
        if g.dtype.kind in "iub" or r.dtype.kind in "iub":
            H.assert_equal(f"{what} {k}", g.astype(np.int64).ravel(), r.astype(np.int64).ravel())
        elif k in exact_float:
            H.assert_equal(f"{what} {k}", g.ravel().view(np.uint32), r.astype(np.float32).ravel().view(np.uint32))
        else:
            H.assert_float_close(f"{what} {k}", g.ravel(), r.astype(np.float32).ravel())
