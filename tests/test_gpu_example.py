"""End-to-end use of the path as a learner would: examples/ppo_on_device.py (sampling with exploration in the kernel,
GAE on the device, torch update) must improve the return of the one-mass oscillator within a few iterations."""
import os.path as osp
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_ppo_on_device_improves_return():
    sys.path.insert(0, osp.join(ROOT, "examples"))
    import ppo_on_device

    hist = ppo_on_device.train(iters=8, n_envs=2048, verbose=False)
    assert len(hist) == 8
    assert hist[-1] > hist[0] + 80.0, hist  # about -437 -> -280 in 8 iterations
    assert all(b > a - 15.0 for a, b in zip(hist, hist[1:])), hist  # no collapse along the way
