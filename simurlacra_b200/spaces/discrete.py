"""DiscreteSpace (pyrado/spaces/discrete.py:39-175): a finite set of elements, one row per element.  Inside the kernels
only the 3-element action set of BallOnBeamDiscSim exists (`SRE_F_DISC_ACT`); this class serves the Python API."""
from typing import Sequence, Union

import numpy as np

from ..utils.exceptions import ShapeErr, TypeErr, ValueErr
from .base import Space


class DiscreteSpace(Space):
    def __init__(self, eles: Union[np.ndarray, list], labels: Sequence[str] = None):
        """
        :param eles: N x D array of all elements, N the number of elements and D the dimension of each
        """
        if isinstance(eles, np.ndarray):
            self.eles = eles if eles.ndim == 2 else eles.reshape(-1, 1)
        elif isinstance(eles, list):
            self.eles = np.array(eles, dtype=np.intc)
            self.eles = self.eles if self.eles.ndim == 2 else self.eles.reshape(-1, 1)
        else:
            raise TypeErr(given=eles, expected_type=[np.ndarray, list])
        self.eles = np.atleast_2d(self.eles)
        self.bound_lo = np.min(self.eles, axis=0)
        self.bound_up = np.max(self.eles, axis=0)
        if labels is not None:
            lab = np.array(labels, dtype=object)
            if not lab.shape == self.shape:
                raise ShapeErr(given=lab, expected_match=self.bound_lo)
            self._labels = lab
        else:
            self._labels = np.empty(self.shape, dtype=object)
            self._labels.fill(None)

    def __str__(self):
        return f"DiscreteSpace(elements={self.eles.tolist()})"

    @property
    def shape(self) -> tuple:
        return self.bound_lo.shape

    @property
    def num_ele(self) -> int:
        return self.eles.shape[0]

    @property
    def flat_dim(self) -> int:
        return self.eles.shape[1]

    @property
    def bound_abs_up(self) -> np.ndarray:
        return np.max(np.abs(self.eles), axis=0)

    @property
    def labels(self) -> np.ndarray:
        return self._labels

    def _members(self) -> tuple:
        return self.eles, self.labels

    def subspace(self, idcs):
        raise NotImplementedError

    def shrink(self, new_lo: np.ndarray, new_up: np.ndarray):
        raise NotImplementedError

    def contains(self, cand: np.ndarray, verbose: bool = False) -> bool:
        cand = np.asarray(cand)
        if not cand.shape == self.shape:
            raise ShapeErr(given=cand, expected_match=self.bound_lo)
        if np.isnan(cand).any():
            raise ValueErr(msg="At least one value is NaN!")
        return bool(np.any(np.isclose(self.eles, cand.astype(self.eles.dtype))))

    def sample_uniform(self, concrete_inf: float = 1e6) -> np.ndarray:
        idx = np.random.randint(self.num_ele, size=1)
        return self.eles[idx, :].flatten()

    def project_to(self, ele: np.ndarray) -> np.ndarray:
        """An element that is (np.isclose) in the space passes unchanged, else the closest element (:119-127)."""
        if not self.contains(ele):
            return self.eles[np.argmin(np.abs(ele - self.eles)), :]
        return ele

    @staticmethod
    def cat(spaces):
        spaces = [s for s in spaces if s is not None]
        eles = []
        for s in spaces:
            if not isinstance(s, DiscreteSpace):
                raise TypeErr(given=s, expected_type=DiscreteSpace)
            eles.extend(s.eles)
        if all(isinstance(e, np.ndarray) for e in eles):
            eles = np.concatenate(eles, axis=0)
        return DiscreteSpace(eles, labels=None)
