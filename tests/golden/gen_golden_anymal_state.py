"""Synthetic Anymal state: lives in the package (isaacgymenvs-ma_b200/synthetic.py); re-exported here for
tests/golden/gen_golden_anymal.py and the GPU parity tests (no reference tree needed)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ige_loader import load_package  # noqa: E402

_syn = load_package().synthetic
CFG, DT = _syn.ANYMAL_CFG, _syn.ANYMAL_DT
advance = _syn.advance_anymal


def make_state(n, seed, reset_prob):
    return _syn.make_anymal(n, seed=seed, reset_prob=reset_prob)
