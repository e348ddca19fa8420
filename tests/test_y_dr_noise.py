"""the observation / action noise lambdas of VecTask.step (vec_task.py:373-374, :399-400).

CPU: the oracle's restatement (oracle/ige_oracle.c::ige_oracle_dr_noise) and the host-side schedule arithmetic
(dr_noise.scheduled_params) against the closures the UNMODIFIED reference builds in VecTask.apply_randomizations
(vec_task.py:644-713) -- same torch draws, bit-exact.  GPU: ige_dr_noise against the oracle (INJECTED: bit-exact;
PHILOX: the same generator, 1e-5 -- device vs libm logf / sincosf) and the task mirror's step() with noise configured
against the reference closure applied to the noise-free step.

(The file name sorts late on purpose: `pytest -x` stops at the first failure, and this newest row must not hide the
parity tests of the rows before it.)
"""
import ctypes as C
import math
import types

import numpy as np
import pytest
import torch

import harness as H
import ref_shim

abi = H.pkg._abi
SPECS = {
    ("gaussian", "additive"): {"range": [0.02, 0.3], "range_correlated": [0.01, 0.05]},
    ("gaussian", "scaling"): {"range": [1.0, 0.1], "range_correlated": [1.0, 0.02], "schedule": "linear", "schedule_steps": 400},
    ("uniform", "additive"): {"range": [-0.2, 0.3], "range_correlated": [-0.05, 0.02], "schedule": "constant", "schedule_steps": 100},
    ("uniform", "scaling"): {"range": [0.8, 1.3], "range_correlated": [0.95, 1.05]},
}
CASES = sorted(SPECS)


def _spec(dist, op):
    return {"distribution": dist, "operation": op, **SPECS[(dist, op)]}


def reference_lambda(name, spec, frame):
    """The reference's own closure: VecTask.apply_randomizations run on a bare object at simulator frame `frame`."""
    vt = ref_shim.ref_module("tasks.base.vec_task")

    class Gym(ref_shim.FakeGym):
        def get_frame_count(self, sim):
            return frame

    obj = types.SimpleNamespace(gym=Gym(record=False), sim=None, first_randomization=True, last_step=-1, last_rand_step=-1,
                                num_envs=4, envs=[], dr_randomizations={}, randomize_buf=torch.zeros(4, dtype=torch.long),
                                reset_buf=torch.ones(4, dtype=torch.long), actor_params_generator=None,
                                original_props={}, extern_actor_params={})
    vt.VecTask.apply_randomizations(obj, {"frequency": 1, name: spec, "actor_params": {}})   # the key is mandatory there (dr_utils.py:212)
    return obj.dr_randomizations[name]


def oracle_noise(x, corr, draw, spec, params, clip):
    from oracle import oracle
    x = np.ascontiguousarray(x, np.float32)
    y, yc = np.empty_like(x), np.empty_like(x)
    corr = np.ascontiguousarray(corr, np.float32).copy()
    draw = np.ascontiguousarray(draw, np.float32)
    a = abi.IgeNoiseArgs()
    a.count = x.size
    a.operation = abi.IGE_NOISE_SCALING if spec["operation"] == "scaling" else abi.IGE_NOISE_ADDITIVE
    a.distribution = abi.IGE_NOISE_UNIFORM if spec["distribution"] == "uniform" else abi.IGE_NOISE_GAUSSIAN
    a.rand_mode = abi.IGE_RAND_INJECTED
    a.corr_scale, a.corr_shift, a.scale, a.shift = params
    a.clip = clip
    a.x, a.y, a.y_clamped = x.ctypes.data, y.ctypes.data, yc.ctypes.data
    a.corr, a.draw = corr.ctypes.data, draw.ctypes.data
    oracle.call("dr_noise", a)
    return y, yc


@pytest.mark.needs_reference
@pytest.mark.parametrize("dist,op", CASES)
@pytest.mark.parametrize("frame", [0, 150, 1000])
def test_oracle_and_schedule_match_reference_closure(dist, op, frame):
    spec = _spec(dist, op)
    ref = reference_lambda("observations", spec, frame)
    params = H.pkg.dr_noise.scheduled_params(spec, frame)
    if dist == "gaussian":
        want = (ref["var_corr"], ref["mu_corr"], ref["var"], ref["mu"])
    else:
        want = (ref["hi_corr"] - ref["lo_corr"], ref["lo_corr"], ref["hi"] - ref["lo"], ref["lo"])
    assert params == want                                   # float64 host arithmetic, identical expression order
    x = torch.from_numpy(np.random.Generator(np.random.PCG64(3)).standard_normal((257, 13)).astype(np.float32) * 3)
    for call in range(2):                                   # 1st call draws corr, 2nd re-uses it (vec_task.py:677-680)
        torch.manual_seed(50 + call)
        want_y = ref["noise_lambda"](x)
        torch.manual_seed(50 + call)
        if call == 0:
            corr = torch.randn_like(x)
        draw = torch.randn_like(x) if dist == "gaussian" else torch.rand_like(x)
        np.testing.assert_array_equal(ref["corr"].numpy(), corr.numpy())
        y, yc = oracle_noise(x.numpy(), corr.numpy(), draw.numpy(), spec, params, 2.5)
        np.testing.assert_array_equal(y.view(np.uint32), want_y.numpy().view(np.uint32))
        np.testing.assert_array_equal(yc.view(np.uint32), torch.clamp(want_y, -2.5, 2.5).numpy().view(np.uint32))


def test_mirror_refuses_simulator_randomisation():
    dr = {"observations": _spec("gaussian", "additive"), "sim_params": {"gravity": {}}}
    with pytest.raises(NotImplementedError, match="sim_params"):
        H.pkg.make(0, "Cartpole", 8, "cpu", dr_randomizations=dr)


def _gpu_noise(x, corr, draw, spec, params, clip, mode, init_corr=0, seed=0, stream_id=0, step=None, in_place=False):
    dev = "cuda:0"
    xd = torch.from_numpy(x).to(dev)
    yd = xd if in_place else torch.full_like(xd, 7.0)
    ycd = torch.full_like(xd, 7.0)
    cd = torch.from_numpy(np.ascontiguousarray(corr, np.float32)).to(dev)
    a = abi.IgeNoiseArgs()
    a.count = x.size
    a.operation = abi.IGE_NOISE_SCALING if spec["operation"] == "scaling" else abi.IGE_NOISE_ADDITIVE
    a.distribution = abi.IGE_NOISE_UNIFORM if spec["distribution"] == "uniform" else abi.IGE_NOISE_GAUSSIAN
    a.rand_mode, a.init_corr = mode, init_corr
    a.corr_scale, a.corr_shift, a.scale, a.shift = params
    a.clip, a.seed, a.stream = clip, seed, stream_id
    a.x, a.y, a.y_clamped, a.corr = xd.data_ptr(), yd.data_ptr(), ycd.data_ptr(), cd.data_ptr()
    dd = None
    if draw is not None:
        dd = torch.from_numpy(np.ascontiguousarray(draw, np.float32)).to(dev)
        a.draw = dd.data_ptr()
    if step is not None:
        sd = torch.tensor([step], dtype=torch.int64, device=dev)
        a.step = sd.data_ptr()
    H.pkg._lib.call("ige_dr_noise", a, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return yd.cpu().numpy(), ycd.cpu().numpy(), cd.cpu().numpy()


@pytest.mark.gpu
@pytest.mark.parametrize("dist,op", CASES)
@pytest.mark.parametrize("shape", [(16384, 60), (1023, 7), (3,), (32768, 108)])
def test_kernel_matches_oracle_injected(dist, op, shape):
    """Bit-exact: the kernel evaluates the closure op by op (no fma), like torch."""
    spec = _spec(dist, op)
    params = H.pkg.dr_noise.scheduled_params(spec, 250)
    rng = np.random.Generator(np.random.PCG64(hash((dist, op, shape)) & 0xffff))
    x = (rng.standard_normal(shape) * 4).astype(np.float32)
    corr = rng.standard_normal(shape).astype(np.float32)
    draw = (rng.standard_normal(shape) if dist == "gaussian" else rng.random(shape)).astype(np.float32)
    want_y, want_yc = oracle_noise(x, corr, draw, spec, params, 5.0)
    for in_place in (False, True):
        y, yc, corr_after = _gpu_noise(x, corr, draw, spec, params, 5.0, abi.IGE_RAND_INJECTED, in_place=in_place)
        np.testing.assert_array_equal(y.view(np.uint32), want_y.view(np.uint32))
        np.testing.assert_array_equal(yc.view(np.uint32), want_yc.view(np.uint32))
        np.testing.assert_array_equal(corr_after, corr)                      # INJECTED never rewrites corr


@pytest.mark.gpu
@pytest.mark.parametrize("dist,op", CASES)
def test_kernel_matches_oracle_philox(dist, op):
    """In-kernel draws: same Philox counters as the oracle's restatement; normals go through device logf / sincosf,
    hence 1e-5 instead of bit-exact.  The first call stores corr, the second re-uses it; moments are sane."""
    from oracle import oracle
    spec = _spec(dist, op)
    params = H.pkg.dr_noise.scheduled_params(spec, 1000)
    n = 16384 * 60 + 3
    x = np.linspace(-3, 3, n, dtype=np.float32)
    zeros = np.zeros(n, np.float32)
    y1, yc1, corr1 = _gpu_noise(x, zeros, None, spec, params, 4.0, abi.IGE_RAND_PHILOX, init_corr=1, seed=9, stream_id=1, step=5)
    assert abs(float(corr1.mean())) < 5e-3 and abs(float(corr1.std()) - 1.0) < 5e-3
    y2, yc2, corr2 = _gpu_noise(x, corr1, None, spec, params, 4.0, abi.IGE_RAND_PHILOX, init_corr=0, seed=9, stream_id=1, step=6)
    np.testing.assert_array_equal(corr2, corr1)
    assert not np.array_equal(y1, y2)                                        # the step word changes the draw
    for step, init, got_y, got_yc in ((5, 1, y1, yc1), (6, 0, y2, yc2)):
        oy, oyc, ocorr = np.empty_like(x), np.empty_like(x), (zeros.copy() if init else corr1.copy())
        stepw = C.c_uint64(step)
        a = abi.IgeNoiseArgs()
        a.count = n
        a.operation = abi.IGE_NOISE_SCALING if op == "scaling" else abi.IGE_NOISE_ADDITIVE
        a.distribution = abi.IGE_NOISE_UNIFORM if dist == "uniform" else abi.IGE_NOISE_GAUSSIAN
        a.rand_mode, a.init_corr = abi.IGE_RAND_PHILOX, init
        a.corr_scale, a.corr_shift, a.scale, a.shift = params
        a.clip, a.seed, a.stream = 4.0, 9, 1
        a.step = C.cast(C.pointer(stepw), C.c_void_p)
        a.x, a.y, a.y_clamped, a.corr = x.ctypes.data, oy.ctypes.data, oyc.ctypes.data, ocorr.ctypes.data
        oracle.call("dr_noise", a)
        if init:
            np.testing.assert_allclose(corr1, ocorr, rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(got_y, oy, rtol=1e-5, atol=2e-5)
        np.testing.assert_allclose(got_yc, oyc, rtol=1e-5, atol=2e-5)
        assert float(np.abs(got_yc).max()) <= 4.0


@pytest.mark.gpu
@pytest.mark.needs_reference
@pytest.mark.parametrize("dist,op", [("gaussian", "additive"), ("uniform", "scaling")])
def test_mirror_step_with_noise_matches_reference_closures(dist, op):
    """Ant mirror, step() with observation AND action noise configured (injected draws): what step() returns equals
    the reference's closures applied around the noise-free step -- actions: noise then clamp (vec_task.py:373-376),
    observations: obs_buf <- noise(obs_buf), obs = clamp(obs_buf) (:399-404)."""
    spec = _spec(dist, op)
    dr = {"frequency": 1, "observations": spec, "actions": spec}
    n = 2048
    st = H.pkg.synthetic.make_state("ant", n, seed=4, reset_prob=0.0)
    mk = lambda extra: H.pkg.factory.from_state("Ant", {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in st.items()},
                                                "cuda:0", rand_mode="injected", seed=1, dr_randomizations=extra)
    plain, noisy = mk(None), mk(dr)
    actions = torch.from_numpy(st["actions"]).float()
    ref_act = reference_lambda("actions", spec, 0)
    ref_obs = reference_lambda("observations", spec, 0)
    torch.manual_seed(77)
    want_actions = torch.clamp(ref_act["noise_lambda"](actions), -plain.clip_actions, plain.clip_actions)
    obs_plain = plain.step(want_actions.to("cuda:0"))[0]["obs"]                 # noise-free step on the noised actions
    want_obs_buf = ref_obs["noise_lambda"](plain.obs_buf.cpu())
    torch.manual_seed(77)
    out = noisy.step(actions.to("cuda:0"))
    torch.cuda.synchronize()
    np.testing.assert_array_equal(noisy.actions.cpu().numpy().view(np.uint32), want_actions.numpy().view(np.uint32))
    np.testing.assert_array_equal(noisy.obs_buf.cpu().numpy().view(np.uint32), want_obs_buf.numpy().view(np.uint32))
    np.testing.assert_array_equal(out[0]["obs"].cpu().numpy().view(np.uint32),
                                  torch.clamp(want_obs_buf, -plain.clip_obs, plain.clip_obs).numpy().view(np.uint32))
    assert not torch.equal(obs_plain, out[0]["obs"])
