"""Polar2DPosVelSpace (pyrado/spaces/polar.py:82-123): samples (r, phi, xd, yd) in a box, returns cartesian
[x, y, xd, yd]; `contains` maps the candidate back to polar coordinates."""
import numpy as np

from .box import BoxSpace


class Polar2DPosVelSpace(BoxSpace):
    def __init__(self, bound_lo, bound_up, shape=None, labels=None):
        super().__init__(bound_lo, bound_up, shape, labels)
        assert self.bound_lo.shape == (4,), "Polar2DPosVelSpace holds [r, phi, xd, yd]"

    def sample_uniform(self, concrete_inf: float = 1e6) -> np.ndarray:
        s = super().sample_uniform(concrete_inf)
        return np.array([s[0] * np.cos(s[1]), s[0] * np.sin(s[1]), s[2], s[3]])

    def contains(self, cand: np.ndarray, verbose: bool = False) -> bool:
        cand = np.asarray(cand, dtype=np.float64)
        polar = np.array([np.hypot(cand[0], cand[1]), np.arctan2(cand[1], cand[0]), cand[2], cand[3]])
        return super().contains(polar, verbose)
