"""Import shim for the *reference* task modules (test infrastructure, container-only).

The reference (/root/reference, read-only, absent on the GPU box) imports the closed-source
``isaacgym`` binary, ``gym`` and hydra at module import time.  None of that is on the hot path
(SURVEY.md section 8c), so this shim installs permissive stub modules and registers empty
``isaacgymenvs`` / ``isaacgymenvs.tasks`` packages whose ``__path__`` points into the reference
tree; the leaf task modules (ant.py, humanoid.py, ...) then import unmodified and their
``post_physics_step`` / ``@torch.jit.script`` functions run on CPU torch.

Only ``tests/golden/gen_golden.py`` (fixture generator) and tests marked ``needs_reference`` use
this.  Nothing in the product package, bench.py or the -m gpu tests may import it.
"""
import importlib
import os
import sys
import types

_STAGED = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")


def _pick_root():
    """IGE_REFERENCE_ROOT, else /root/reference (build container), else the unmodified files staged under the
    git-ignored oracle/_ref/ by tools/vendor_reference.py (this is what travels to the GPU box)."""
    env = os.environ.get("IGE_REFERENCE_ROOT")
    if env:
        return env
    for cand in ("/root/reference", _STAGED):
        if os.path.isdir(os.path.join(cand, "isaacgymenvs", "tasks")):
            return cand
    return "/root/reference"


REFERENCE_ROOT = _pick_root()


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "isaacgymenvs", "tasks"))


class _Permissive(types.ModuleType):
    """Module whose unknown attributes resolve to inert callables / constants."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        val = _Anything(name)
        setattr(self, name, val)
        return val


class _Anything:
    def __init__(self, name="x"):
        self._name = name

    def __call__(self, *a, **k):
        return _Anything(self._name + "()")

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything(self._name + "." + name)


class Unwrapped:
    """What the stub ``gymtorch.unwrap_tensor`` returns: a marker around the torch tensor.  The real
    ``gymapi.Gym`` methods take these descriptors and reject torch tensors; FakeGym does the same, so a
    binding that forgets ``unwrap_tensor`` fails here too instead of only under Isaac Gym."""

    def __init__(self, tensor):
        import torch
        if not isinstance(tensor, torch.Tensor):
            raise TypeError(f"gymtorch.unwrap_tensor expects a torch tensor, got {type(tensor).__name__}")
        self.tensor = tensor


def install():
    """Idempotently install the stubs; returns the dict of stub modules."""
    if "isaacgym" in sys.modules and getattr(sys.modules["isaacgym"], "_ige_stub", False):
        return
    import numpy as np
    if not hasattr(np, "Inf"):
        np.Inf = np.inf  # numpy 2 removed the alias used at vec_task.py:108-117

    ig = _Permissive("isaacgym")
    ig._ige_stub = True
    ig.__path__ = []
    gymapi = _Permissive("isaacgym.gymapi")
    gymtorch = _Permissive("isaacgym.gymtorch")
    gymutil = _Permissive("isaacgym.gymutil")
    gymtorch.__all__ = []            # ant.py:35 does `from isaacgym.gymtorch import *`
    gymtorch.wrap_tensor = lambda t: t
    gymtorch.unwrap_tensor = Unwrapped   # FakeGym(strict) refuses bare torch tensors, as the real gym API does
    ig.gymapi, ig.gymtorch, ig.gymutil = gymapi, gymtorch, gymutil
    sys.modules.update({"isaacgym": ig, "isaacgym.gymapi": gymapi,
                        "isaacgym.gymtorch": gymtorch, "isaacgym.gymutil": gymutil})

    gym = _Permissive("gym")
    gym.__path__ = []
    spaces = _Permissive("gym.spaces")

    class Box:  # gym.spaces.Box stand-in (vec_task.py:108-117)
        def __init__(self, low, high, *a, **k):
            self.low, self.high = low, high
            self.shape = getattr(low, "shape", None)

    spaces.Box = Box
    gym.spaces = spaces
    sys.modules.update({"gym": gym, "gym.spaces": spaces})

    for name, rel in (("isaacgymenvs", "isaacgymenvs"),
                      ("isaacgymenvs.tasks", "isaacgymenvs/tasks"),
                      ("isaacgymenvs.tasks.base", "isaacgymenvs/tasks/base"),
                      ("isaacgymenvs.utils", "isaacgymenvs/utils"),
                      ("isaacgymenvs.learning", "isaacgymenvs/learning")):
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, rel)]
        sys.modules[name] = pkg


def ref_module(name: str):
    """Import e.g. ``ref_module('tasks.ant')`` -> isaacgymenvs.tasks.ant from the reference."""
    install()
    return importlib.import_module("isaacgymenvs." + name)


class FakeGym:
    """Stand-in for ``gymapi.acquire_gym()``: refresh_* are no-ops (the synthetic tensors already
    hold post-simulate state), set_*_indexed calls are *recorded* and never mutate anything
    (SURVEY.md section 8c)."""

    def __init__(self, record=True):
        self.calls = []
        self.record = record          # False: set_* calls are dropped (timed baseline runs)
        # True: a bare torch tensor in a set_* call raises, as under the real gym API.  Opt-in, because some
        # reference tasks pass the descriptor they got from gym.acquire_* (ball_balance.py:397), which the
        # harness also models as a torch tensor.
        self.strict = False

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        if name.startswith("refresh_"):
            return lambda *a, **k: None
        if not self.record:
            return lambda *a, **k: None

        def rec(sim, *args):
            import torch
            saved = []
            for a in args:
                if self.strict and isinstance(a, torch.Tensor):
                    raise TypeError(f"gym.{name}: got a torch tensor; pass gymtorch.unwrap_tensor(t) (ant.py:265-271)")
                if isinstance(a, Unwrapped):
                    a = a.tensor
                if isinstance(a, torch.Tensor) and a.dtype == torch.int32:
                    saved.append(a.clone())
                elif isinstance(a, torch.Tensor):
                    saved.append(("tensor", a.data_ptr(), tuple(a.shape)))
                else:
                    saved.append(a)
            self.calls.append((name, saved))
        return rec
