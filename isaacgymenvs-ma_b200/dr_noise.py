"""Observation / action noise of ``VecTask.step`` on top of ``ige_dr_noise``.

``NoiseLambda`` stands where the reference keeps ``dr_randomizations[name]['noise_lambda']`` (vec_task.py:373-374
actions, :399-400 observations): a callable on the tensor, holding the persistent correlated sample ``corr`` and
the four numbers apply_randomizations derives from the config and the schedule (vec_task.py:644-713).  The device
work -- noise, operation and the clamp of step() that follows -- is one launch of the CUDA kernel; there is no torch
fallback.  Only these two keys of the randomisation config belong to the hot path; simulation / actor parameter
randomisation is simulator set-up (OUT OF SCOPE) and is refused by the task mirror.
"""
import math

import torch

from . import _abi, _lib

HOT_PATH_KEYS = ("observations", "actions")


def scheduled_params(spec, last_step):
    """(corr_scale, corr_shift, scale, shift) as float64 host numbers -- vec_task.py:644-673 (schedule),
    :660-673 gaussian, :689-701 uniform.  ``spec`` is the reference's config entry: distribution, operation, range,
    optional range_correlated / schedule / schedule_steps."""
    dist, op = spec["distribution"], spec["operation"]
    if dist not in ("gaussian", "uniform") or op not in ("additive", "scaling"):
        raise ValueError(f"noise spec: distribution={dist!r} operation={op!r} not supported")
    sched, steps = spec.get("schedule"), spec.get("schedule_steps")
    if sched == "linear":
        k = 1.0 / steps * min(last_step, steps)
    elif sched == "constant":
        k = 0 if last_step < steps else 1
    else:
        k = 1
    a, b = spec["range"]
    ac, bc = spec.get("range_correlated", [0.0, 0.0])
    if dist == "gaussian":                      # (mu, var), (mu_corr, var_corr)
        mu, var, mu_c, var_c = a, b, ac, bc
        if op == "additive":
            mu, var, mu_c, var_c = mu * k, var * k, mu_c * k, var_c * k
        else:
            var, mu = var * k, mu * k + 1.0 * (1.0 - k)
            var_c, mu_c = var_c * k, mu_c * k + 1.0 * (1.0 - k)
        return var_c, mu_c, var, mu
    lo, hi, lo_c, hi_c = a, b, ac, bc
    if op == "additive":
        lo, hi, lo_c, hi_c = lo * k, hi * k, lo_c * k, hi_c * k
    else:
        lo, hi = lo * k + 1.0 * (1.0 - k), hi * k + 1.0 * (1.0 - k)
        lo_c, hi_c = lo_c * k + 1.0 * (1.0 - k), hi_c * k + 1.0 * (1.0 - k)
    return hi_c - lo_c, lo_c, hi - lo, lo


class NoiseLambda:
    """``y = noise(tensor)`` as the reference's closure computes it, plus ``clamp(y, +-clip)``.

    ``rand_mode="philox"``: draws made in the kernel (key = seed, counter = element, tensor id, step word).
    ``rand_mode="injected"``: ``__call__`` takes ``draw=`` (the reference's ``randn_like`` / ``rand_like``) and, on the
    first call, ``corr=`` (its first ``randn_like``) -- the parity mode."""

    def __init__(self, spec, device, clip=math.inf, tensor_id=0, seed=0, step=None, rand_mode="philox"):
        self.spec, self.device = dict(spec), torch.device(device)
        if self.device.type != "cuda":
            raise _lib.IgeError("ige_dr_noise needs CUDA tensors; there is no CPU path")
        self.clip, self.tensor_id, self.seed, self.step = float(clip), int(tensor_id), int(seed), step
        self.rand_mode = rand_mode
        self.corr = None                      # params['corr'] (vec_task.py:677-680)
        self.calls = 0
        self.set_schedule(0)

    def set_schedule(self, last_step):
        """apply_randomizations recomputes the numbers every time it runs (every `frequency` steps)."""
        self.params = scheduled_params(self.spec, last_step)

    def __call__(self, tensor, out=None, out_clamped=None, draw=None, corr=None):
        if not (tensor.is_cuda and tensor.is_contiguous() and tensor.dtype == torch.float32):
            raise _lib.IgeError("ige_dr_noise: tensor must be a contiguous float32 CUDA tensor")
        out = torch.empty_like(tensor) if out is None else out
        a = _abi.IgeNoiseArgs()
        a.count = tensor.numel()
        a.operation = _abi.IGE_NOISE_SCALING if self.spec["operation"] == "scaling" else _abi.IGE_NOISE_ADDITIVE
        a.distribution = _abi.IGE_NOISE_UNIFORM if self.spec["distribution"] == "uniform" else _abi.IGE_NOISE_GAUSSIAN
        a.corr_scale, a.corr_shift, a.scale, a.shift = self.params
        a.clip, a.stream, a.seed = self.clip, self.tensor_id, self.seed
        a.step = self.step.data_ptr() if self.step is not None else None
        first = self.corr is None or self.corr.shape != tensor.shape
        if self.rand_mode == "philox":
            a.rand_mode, a.init_corr = _abi.IGE_RAND_PHILOX, int(first)
            if first:
                self.corr = torch.empty_like(tensor)
        else:
            a.rand_mode, a.init_corr = _abi.IGE_RAND_INJECTED, 0
            if first:
                if corr is None:
                    raise ValueError("injected mode: the first call needs corr= (the reference's first randn_like)")
                self.corr = corr.to(self.device, torch.float32).contiguous()
            if draw is None:
                raise ValueError("injected mode: draw= is required")
            draw = draw.to(self.device, torch.float32).contiguous()
            a.draw = draw.data_ptr()
        a.x, a.y, a.corr = tensor.data_ptr(), out.data_ptr(), self.corr.data_ptr()
        a.y_clamped = out_clamped.data_ptr() if out_clamped is not None else None
        _lib.call("ige_dr_noise", a, stream=torch.cuda.current_stream(self.device).cuda_stream)
        self._keep = draw                     # alive until the stream has consumed it
        self.calls += 1
        return out
