"""CompoundSpace (pyrado/spaces/compound.py:38-90): a union of spaces; sampling picks one uniformly at random."""
from typing import Sequence

import numpy as np

from .base import Space


class CompoundSpace(Space):
    def __init__(self, spaces: Sequence[Space]):
        self._spaces = list(spaces)
        self.bound_lo = np.min(np.stack([s.bound_lo for s in self._spaces]), axis=0)
        self.bound_up = np.max(np.stack([s.bound_up for s in self._spaces]), axis=0)

    @property
    def shape(self):
        return self._spaces[0].shape

    def _members(self):
        return tuple(m for s in self._spaces for m in s._members())

    def project_to(self, ele: np.ndarray):
        raise NotImplementedError

    def subspace(self, idcs):
        return CompoundSpace([s.subspace(idcs) for s in self._spaces])

    def shrink(self, new_lo, new_up):
        raise NotImplementedError("Cannot shrink a compound space")

    def contains(self, cand: np.ndarray, verbose: bool = False) -> bool:
        return any(s.contains(cand) for s in self._spaces)

    def sample_uniform(self, concrete_inf: float = 1e6) -> np.ndarray:
        return self._spaces[np.random.randint(len(self._spaces))].sample_uniform(concrete_inf)
