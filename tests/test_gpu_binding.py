"""The drop-in claim, executed: ``libige_b200.so`` bound under the REFERENCE's own, unmodified task classes.

``integration.bind(isaacgymenvs.tasks.ant.Ant)`` replaces the one virtual the hot path plugs into
(``post_physics_step``, vec_task.py:358-360); the reference's own ``VecTask.step`` / ``pre_physics_step`` then run
on CUDA tensors (FakeGym in place of the simulator) and every buffer ``step()`` exposes is compared
with the same class, un-patched, on CPU torch: ids / flags / counters bit-exact, floats within 1e-5 relative.

Needs the reference tree: /root/reference in the build container, the staged copy under oracle/_ref/ (written by
``__graft_entry__.build()`` via tools/vendor_reference.py) on the GPU box.
"""
import numpy as np
import pytest
import torch

import harness as H
import ref_harness
import ref_shim

pytestmark = [pytest.mark.gpu, pytest.mark.needs_reference]
syn = H.pkg.synthetic
DEV = "cuda:0"


def _to_device(obj, n, d):
    """Move every tensor attribute of a reference task object to the GPU and rebuild its strided views."""
    for k, v in list(vars(obj).items()):
        if isinstance(v, torch.Tensor):
            setattr(obj, k, v.to(DEV))
    obj.dof_pos = obj.dof_state.view(n, d, 2)[..., 0]        # ant.py:94-95
    obj.dof_vel = obj.dof_state.view(n, d, 2)[..., 1]
    obj.device = DEV
    obj.rl_device = DEV
    return obj


def _build(task, cls, st, cfg):
    obj = ref_harness._build_locomotion(cls, {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in st.items()}, cfg)
    if task == "ant":
        obj.joint_gears = torch.full((cfg["num_dof"],), 15.0)
    else:
        obj.angular_velocity_scale = cfg["angular_velocity_scale"]
        obj.dof_force_tensor = ref_harness._t(st["dof_force_tensor"])
        obj.motor_efforts = ref_harness._t(st["motor_efforts"])
        obj.max_motor_effort = float(obj.motor_efforts.max())
    return obj


def _calls(obj):
    return [(name, [a.cpu().numpy() for a in args if isinstance(a, torch.Tensor) and a.dtype == torch.int32])
            for name, args in obj.gym.calls if "indexed" in name]


@pytest.mark.parametrize("task,n", [("ant", 4096), ("humanoid", 2048), ("ant", 1000)])
def test_fused_post_physics_under_reference_step(task, n):
    cfg = {"ant": syn.ANT_CFG, "humanoid": syn.HUMANOID_CFG}[task]
    d = cfg["num_dof"]
    mod = ref_shim.ref_module("tasks." + task)
    ref_cls = getattr(mod, task.capitalize())
    fused_cls = H.pkg.integration.bind(ref_cls)
    assert fused_cls.__mro__[2] is ref_cls and fused_cls.step is ref_cls.step        # VecTask.step is the reference's
    assert fused_cls.pre_physics_step is ref_cls.pre_physics_step

    st = syn.make_state(task, n, seed=11, reset_prob=0.05)
    ref = _build(task, ref_cls, st, cfg)                                              # un-patched, CPU
    fused = _to_device(_build(task, fused_cls, st, cfg), n, d)                        # patched, CUDA
    ref.gym.strict = fused.gym.strict = True      # bare torch tensors in gym.set_* calls raise (ant.py:265-271 wraps them)
    fused.ige_rand_mode = "injected"
    # the reference draws its reset noise from torch's CPU generator: draw the same stream, then move it over
    fused.ige_draw_reset_uniforms = lambda k: tuple(torch.rand((k, d)).to(DEV) for _ in range(2))
    fused.ige_setup()

    rng = np.random.Generator(np.random.PCG64(5))
    for s in range(3):
        if s > 0:
            for name in ref_harness.OUTPUTS[task]:
                st[name] = getattr(ref, name).detach().numpy().copy()
            syn.advance_physics(task, st, rng)
        for name in ref_harness.SIM_INPUTS[task]:
            if name != "actions":
                getattr(ref, name).copy_(ref_harness._t(st[name]))
                getattr(fused, name).copy_(ref_harness._t(st[name]).to(DEV))
        actions = ref_harness._t(st["actions"])
        ref.gym.calls.clear()
        fused.gym.calls.clear()
        torch.manual_seed(100 + s)
        obs_r, rew_r, reset_r, extras_r = ref.step(actions)                          # the reference, end to end
        torch.manual_seed(100 + s)
        obs_f, rew_f, reset_f, extras_f = fused.step(actions.to(DEV))                # same step(), fused virtual
        torch.cuda.synchronize()
        what = f"{task} step {s}"
        H.assert_equal(f"{what} reset", reset_f.cpu().numpy(), reset_r.numpy())
        H.assert_equal(f"{what} progress_buf", fused.progress_buf.cpu().numpy(), ref.progress_buf.numpy())
        H.assert_equal(f"{what} randomize_buf", fused.randomize_buf.cpu().numpy(), ref.randomize_buf.numpy())
        H.assert_equal(f"{what} time_outs", extras_f["time_outs"].cpu().numpy(), extras_r["time_outs"].numpy())
        H.assert_float_close(f"{what} obs", obs_f["obs"].cpu().numpy(), obs_r["obs"].numpy())
        H.assert_float_close(f"{what} rew", rew_f.cpu().numpy(), rew_r.numpy())
        for name in ("potentials", "prev_potentials", "dof_state"):                   # reset rows: bit-exact
            H.assert_equal(f"{what} {name}", getattr(fused, name).cpu().numpy().view(np.uint32),
                           getattr(ref, name).numpy().view(np.uint32))
        for name in ("up_vec", "heading_vec"):
            H.assert_float_close(f"{what} {name}", getattr(fused, name).cpu().numpy(), getattr(ref, name).numpy())
        if task == "ant":
            H.assert_equal(f"{what} true_objective", extras_f["true_objective"].cpu().numpy().view(np.uint32),
                           extras_r["true_objective"].numpy().view(np.uint32))
        # the simulator sees the same indexed calls with the same int32 id lists (ant.py:263-271)
        cr, cf = _calls(ref), _calls(fused)
        assert [c[0] for c in cr] == [c[0] for c in cf], what
        for (_, a), (_, b) in zip(cr, cf):
            assert len(a) == len(b) == 1 and np.array_equal(a[0], b[0]), what
    assert int((ref.reset_buf != 0).sum()) > 0 and len(_calls(ref)) > 0               # resets did happen
