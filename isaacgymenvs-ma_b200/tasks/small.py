"""BallBalance, Quadcopter and Ingenuity mirrors (reference: isaacgymenvs/tasks/ball_balance.py, quadcopter.py,
ingenuity.py) -- SURVEY.md 8f row 2.

These tasks run ``reset_idx`` at the top of ``pre_physics_step`` and compute observations / rewards in
``post_physics_step``; each half is ONE C-ABI call (``ige_<task>_pre_physics`` / ``ige_<task>_post_physics``).  The
pre-physics kernel emits the ascending reset id list the reference hands to ``gym.set_*_tensor_indexed``.
Attribute names and tensor layouts follow the reference classes.  No CPU fallback."""
import math

import torch

from .. import _abi, _lib
from .base import VecTaskB200, _ptr

# torch.rand draws of the reset path, in the reference's call order: columns per draw
DRAW_COLS = {
    "ball_balance": (1, 1, 1, 1, 1),          # ball_balance.py:373-383
    "quadcopter": (1, 1, 1, 8),               # quadcopter.py:289-294
    "ingenuity": (2, 1, 2, 1, 1, 1, 1),       # ingenuity.py:286-287 (x2: pre_physics_step :327, reset_idx :304), :309-311
}
# tensors of each task: name -> trailing shape (leading dim = num_envs); S = owned by the simulator
TENSORS = {
    "ball_balance": dict(root_states=(2, 13), dof_states=(6, 2), sensor_tensor=(3, 6), initial_root_states=(2, 13),
                         initial_dof_states=(6, 2), dof_position_targets=(6,), actions=(3,)),
    "quadcopter": dict(root_states=(13,), dof_states=(8, 2), initial_root_states=(13,), initial_dof_states=(8, 2),
                       dof_position_targets=(8,), thrusts=(4,), forces=(9, 3), actions=(12,)),
    "ingenuity": dict(root_tensor=(2, 13), dof_states=(4, 2), initial_root_states=(13,), target_root_positions=(3,),
                      thrusts=(2, 3), forces=(6, 3), actions=(6,)),
}
SIM_TENSORS = {"ball_balance": ("root_states", "dof_states", "sensor_tensor"), "quadcopter": ("root_states", "dof_states"),
               "ingenuity": ("root_tensor", "dof_states")}
ARGS = {"ball_balance": _abi.IgeBallBalanceArgs, "quadcopter": _abi.IgeQuadcopterArgs, "ingenuity": _abi.IgeIngenuityArgs}
NUM_OBS = {"ball_balance": 24, "quadcopter": 21, "ingenuity": 13}


def fill_scalars(task, a, cfg, num_envs):
    """cfg -> the scalar fields of the task's Args struct.  Products of Python floats are formed in double precision
    and rounded once, as the reference's expressions do before they meet a tensor."""
    a.num_envs = num_envs
    a.max_episode_length = float(cfg["max_episode_length"])
    a.clip_obs = float(cfg.get("clip_obs", math.inf))
    dt = float(cfg["dt"])
    if task == "ball_balance":
        a.ball_radius = float(cfg["ball_radius"])
        a.target_step = dt * float(cfg["action_speed_scale"])                    # ball_balance.py:421
        _abi.fill_array(a.dof_lower, cfg["dof_lower"])
        _abi.fill_array(a.dof_upper, cfg["dof_upper"])
    elif task == "quadcopter":
        a.num_bodies = int(cfg["num_bodies"])
        a.dof_target_step = dt * (8 * math.pi)                                   # quadcopter.py:309-310
        a.thrust_step = dt * 200                                                 # quadcopter.py:313-314
        _abi.fill_array(a.dof_lower, cfg["dof_lower"])
        _abi.fill_array(a.dof_upper, cfg["dof_upper"])
        _abi.fill_array(a.thrust_lower, [0.0] * 4)                               # quadcopter.py:88-90
        _abi.fill_array(a.thrust_upper, [float(cfg["max_thrust"])] * 4)
        _abi.fill_array(a.rotor_bodies, cfg["rotor_bodies"])
    else:
        a.num_bodies = int(cfg["num_bodies"])
        a.dt = dt
        a.thrust_upper_limit = float(cfg["thrust_upper_limit"])
        a.thrust_lateral_component = float(cfg["thrust_lateral_component"])
    return a


class PreResetTask(VecTaskB200):
    """Common mirror of the three classes; subclasses only name the task."""
    TASK = None
    FUSED_ACTION_CLAMP = False

    def __init__(self, cfg, sim, device, **kw):
        cfg = dict(cfg)
        cfg.setdefault("num_obs", NUM_OBS[self.TASK])
        cfg.setdefault("num_actions", TENSORS[self.TASK]["actions"][0])
        super().__init__(cfg, sim, device, **kw)
        n, dev = self.num_envs, self.device
        self.max_episode_length = int(cfg["max_episode_length"])
        self.dt = float(cfg["dt"])
        for name, shape in TENSORS[self.TASK].items():
            t = sim.tensors.get(name)
            if t is None:
                t = torch.zeros((n,) + shape, device=dev)
            setattr(self, name, t.view((n,) + shape))
        if self.TASK == "ingenuity":
            self.target_rank = torch.zeros(n, dtype=torch.int32, device=dev)
            self.actor_dirty = torch.zeros(n * 2, dtype=torch.int64, device=dev)
            self.root_states = self.root_tensor[:, 0]                           # ingenuity.py:70, :80
            self.marker_states = self.root_tensor[:, 1]
        self._u = [None] * len(DRAW_COLS[self.TASK])
        self._args = None

    def adopt_state(self, st):
        for name in ("reset_buf", "progress_buf", *TENSORS[self.TASK]):
            if name in st and name not in SIM_TENSORS[self.TASK]:
                getattr(self, name).copy_(torch.from_numpy(st[name]).to(self.device).view_as(getattr(self, name)))
        self.actions.copy_(torch.clamp(self.actions, -self.clip_actions, self.clip_actions))   # vec_task.py:373

    def _on_remap(self):
        self._args = None

    def abi_struct(self):
        """The task's Args struct; pointers are stable, so it is built once (no host sync: safe inside a capture)."""
        a = self._args
        if a is None:
            a = self._args = fill_scalars(self.TASK, ARGS[self.TASK](), self.cfg, self.num_envs)
            for name in TENSORS[self.TASK]:
                setattr(a, name, _ptr(getattr(self, name)))
            a.obs_buf, a.rew_buf = _ptr(self.obs_buf), _ptr(self.rew_buf)
            a.reset_buf, a.progress_buf = _ptr(self.reset_buf), _ptr(self.progress_buf)
            a.obs_clamped = _ptr(self.obs_clamped) if self.fuse_epilogue else None
            a.timeout_buf = _ptr(self.timeout_bool) if self.fuse_epilogue else None
            a.list = self._list_struct()
            if self.TASK == "ingenuity":
                a.actor_dirty = _ptr(self.actor_dirty)
        if self.rand_mode == "philox":
            a.rand.mode, a.rand.seed, a.rand.step = _abi.IGE_RAND_PHILOX, self.seed, _ptr(self.rng_step)
        else:
            a.rand.mode, a.rand.rank = _abi.IGE_RAND_INJECTED, _ptr(self.reset_rank)
            for i, u in enumerate(self._u):
                a.rand.u[i] = u.data_ptr() if u is not None and u.numel() else None
            if self.TASK == "ingenuity":
                a.target_rank = _ptr(self.target_rank)
        return a

    def draw_reset_uniforms(self, k, kt=0):
        """INJECTED mode: the torch.rand calls of the reset path, reference order and shapes (k reset envs; Ingenuity:
        kt envs with progress % 500 == 0 draw first, ingenuity.py:325-327)."""
        cols = DRAW_COLS[self.TASK]
        out = []
        for i, c in enumerate(cols):
            rows = kt if (self.TASK == "ingenuity" and i < 2) else k
            out.append(torch.rand((rows, c), device=self.device) if rows > 0 else None)
        return out

    def pre_physics_step(self, actions):
        self.actions.copy_(actions.to(self.device).view_as(self.actions))
        self._pre_launch()

    def _pre_launch(self):
        if self.rand_mode != "philox":
            self.compact_resets()                       # reset_rank (row of each reset env in the injected draws)
            kt = 0
            if self.TASK == "ingenuity":
                hit = (self.progress_buf % 500 == 0)
                self.target_rank.copy_((torch.cumsum(hit, 0) - hit.long()).int())
                kt = int(hit.sum())
            k = int(self.reset_count.item())            # the reference's own host sync: len(reset_env_ids)
            if all(u is None for u in self._u):
                self._u = self.draw_reset_uniforms(k, kt)
        _lib.call(f"ige_{self.TASK}_pre_physics", self.abi_struct(), stream=self._stream())
        self._u = [None] * len(self._u)
        self._apply_sim_calls()

    def _apply_sim_calls(self):
        ids, cnt = self.reset_env_ids, self.reset_count
        if self.TASK == "ball_balance":                 # ball_balance.py:393-399, :427
            self.sim.set_indexed("set_actor_root_state_tensor_indexed", self.root_states, ids, cnt)
            self.sim.set_indexed("set_dof_state_tensor_indexed", self.dof_states, ids, cnt)
            self.sim.set_indexed("set_dof_position_target_tensor", self.dof_position_targets, None, None)
        elif self.TASK == "quadcopter":                 # quadcopter.py:292-296, :328-329
            self.sim.set_indexed("set_actor_root_state_tensor_indexed", self.root_states, ids, cnt)
            self.sim.set_indexed("set_dof_state_tensor_indexed", self.dof_states, ids, cnt)
            self.sim.set_indexed("set_dof_position_target_tensor", self.dof_position_targets, None, None)
            self.sim.set_indexed("apply_rigid_body_force_tensors", self.forces, None, None)
        else:                                           # ingenuity.py:314, :341-343, :364 (actor_dirty = unique() flags)
            self.sim.set_indexed("set_dof_state_tensor_indexed", self.dof_states, ids, cnt)
            self.sim.set_indexed("set_actor_root_state_tensor_indexed", self.root_tensor, self.actor_dirty, None)
            self.sim.set_indexed("apply_rigid_body_force_tensors", self.forces, None, None)

    def post_physics_step(self):
        _lib.call(f"ige_{self.TASK}_post_physics", self.abi_struct(), stream=self._stream())

    def hot_step(self):
        """Both halves belong to the path: reset (pre-physics position) + obs / reward / done (post-physics)."""
        self._pre_launch()
        self.post_physics_step()


class BallBalance(PreResetTask):
    TASK = "ball_balance"


class Quadcopter(PreResetTask):
    TASK = "quadcopter"


class Ingenuity(PreResetTask):
    TASK = "ingenuity"
