"""the torch_jit_utils helpers on their own.  CPU: the oracle's restatements against the reference's
own functions (utils/torch_jit_utils.py, imported unmodified).  GPU: the __device__ versions against the oracle."""
import math

import numpy as np
import pytest
import torch

import harness as H

abi = H.pkg._abi


def _inputs(n, seed=0):
    g = np.random.Generator(np.random.PCG64(seed))
    q = g.standard_normal((n, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    q2 = g.standard_normal((n, 4)).astype(np.float32)
    q2 /= np.linalg.norm(q2, axis=1, keepdims=True)
    v = (g.standard_normal((n, 3)) * np.array([4.0, 1.0, 1.0])).astype(np.float32)
    v[:, 2] = v[:, 1] + np.abs(v[:, 2]) + 0.1          # unscale: upper > lower
    v[: n // 16] *= 1e-11                               # normalize(): the 1e-9 clamp branch
    v[: n // 16, 2] = v[: n // 16, 1] + 0.5
    return q, q2, v


def _probe(device, q, q2, v):
    n = q.shape[0]
    t = [torch.from_numpy(np.ascontiguousarray(x)).to(device) for x in (q, q2, v)]
    out = torch.zeros((n, abi.IGE_PROBE_COLS), device=device)
    a = abi.IgeHelperProbeArgs()
    a.n, a.q, a.q2, a.v, a.out = n, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), out.data_ptr()
    if device == "cpu":
        H.oracle.call("debug_helper_probe", a)
    else:
        H.pkg._lib.call("ige_debug_helper_probe", a, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.needs_reference
def test_oracle_helpers_match_reference_torch_jit_utils():
    import ref_shim
    tj = ref_shim.ref_module("utils.torch_jit_utils")
    q, q2, v = _inputs(4096)
    got = _probe("cpu", q, q2, v)
    tq, tq2, tv = torch.from_numpy(q), torch.from_numpy(q2), torch.from_numpy(v)
    roll, _pitch, yaw = tj.get_euler_xyz(tq)
    want = {
        (0, 4): tj.quat_mul(tq, tq2), (4, 7): tj.quat_rotate(tq, tv), (7, 10): tj.quat_rotate_inverse(tq, tv),
        (10, 13): tj.normalize(tv), (13, 14): roll[:, None], (14, 15): yaw[:, None],
        (15, 16): tj.normalize_angle(tv[:, 0])[:, None], (16, 20): tj.quat_conjugate(tq),
        (20, 21): torch.norm(tv, p=2, dim=-1)[:, None],
        (21, 22): torch.bmm(tq[:, :3].view(-1, 1, 3), tv.view(-1, 3, 1)).view(-1, 1),
        (22, 23): tj.unscale(tv[:, 0], tv[:, 1], tv[:, 2])[:, None], (23, 24): (tv[:, 0] % (2 * np.pi))[:, None],
    }
    # pure IEEE arithmetic in one fixed order: bit-exact.  quat_rotate / quat_rotate_inverse: `a + b + c` of three products
    # whose last ulp depends on how TorchScript's CPU fuser contracts them -- 1e-6, and > 90 % of the words identical
    exact = {(0, 4), (10, 13), (16, 20), (20, 21), (21, 22), (22, 23), (23, 24)}
    for (a, b), w in want.items():
        w = w.numpy().astype(np.float32)
        if (a, b) in exact:
            H.assert_equal(f"helper columns {a}:{b}", got[:, a:b].view(np.uint32), w.view(np.uint32))
        elif (a, b) in {(4, 7), (7, 10)}:
            H.assert_float_close(f"helper columns {a}:{b}", got[:, a:b], w, rtol=1e-6)
            assert H.bitexact_fraction(got[:, a:b], w) > 0.9
        else:       # atan2 / sin / cos: libm vs Sleef differ by ulps; angles near the 0 / 2 pi seam may wrap
            d = np.abs(got[:, a:b].astype(np.float64) - w)
            d = np.minimum(d, np.abs(d - 2 * math.pi))
            assert (d <= 1e-5 * np.maximum(1.0, np.abs(w))).all(), (a, b, float(d.max()))


@pytest.mark.gpu
def test_device_helpers_match_oracle():
    q, q2, v = _inputs(100003, seed=1)
    got, want = _probe("cuda:0", q, q2, v), _probe("cpu", q, q2, v)
    trig = [13, 14, 15]
    cols = [c for c in range(abi.IGE_PROBE_COLS) if c not in trig]
    H.assert_equal("arithmetic helper columns (bit-exact)", got[:, cols].view(np.uint32), want[:, cols].view(np.uint32))
    d = np.abs(got[:, trig].astype(np.float64) - want[:, trig])
    d = np.minimum(d, np.abs(d - 2 * math.pi))
    assert (d <= 1e-5 * np.maximum(1.0, np.abs(want[:, trig]))).all(), float(d.max())
