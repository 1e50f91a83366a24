"""Space interface (pyrado/spaces/base.py:43-205).  Spaces are host-side descriptors: inside the kernels the same
geometry is evaluated per environment (bounds in registers); these classes serve the Python API (`env.obs_space`,
`spec`, single-element `contains` / `project_to` / `sample_uniform`)."""
from abc import ABC, abstractmethod
from copy import deepcopy
from typing import Tuple, Union

import numpy as np


class Space(ABC):
    bound_lo: np.ndarray
    bound_up: np.ndarray

    @property
    def bounds(self) -> Tuple[np.ndarray, np.ndarray]:
        return self.bound_lo, self.bound_up

    @property
    def bound_abs_up(self) -> np.ndarray:
        return np.max(np.stack([np.abs(self.bound_lo), np.abs(self.bound_up)], axis=0), axis=0)

    @property
    def ele_dim(self) -> int:
        return int(np.prod(self.bound_lo.shape))

    @property
    @abstractmethod
    def shape(self) -> tuple:
        raise NotImplementedError

    @property
    def flat_dim(self) -> int:
        return int(np.prod(self.shape))

    @property
    def labels(self):
        return None

    @abstractmethod
    def _members(self) -> tuple:
        raise NotImplementedError

    def __eq__(self, other):
        if type(self) is not type(other):
            return False
        return all(np.array_equal(a, b) for a, b in zip(self._members(), other._members()))

    def __str__(self):
        return f"{type(self).__name__}(lo={self.bound_lo}, up={self.bound_up})"

    def copy(self):
        return deepcopy(self)

    def create_mask(self, *idcs) -> np.ndarray:
        """Boolean mask with True at the given indices; an index is a number or one of the space's labels
        (spaces/base.py:103-136)."""
        from ..utils.exceptions import TypeErr, ValueErr

        mask = np.zeros(self.shape, dtype=bool)
        if len(idcs) == 1 and not isinstance(idcs[0], str) and hasattr(idcs[0], "__iter__"):
            idcs = idcs[0]
        for idx in idcs:
            if isinstance(idx, str):
                if self.labels is None:
                    raise TypeErr(msg="The space must be labeled to use label-based indexing!")
                hits = np.argwhere(np.asarray(self.labels) == idx)
                if len(hits) == 0:
                    raise ValueErr(msg=f"Label {idx} not found in {self}")
                idx = tuple(hits[0])
            if np.all(mask[idx]):
                raise ValueErr(msg=f"Duplicate index {idx}")
            mask[idx] = True
        return mask

    @abstractmethod
    def subspace(self, idcs: Union[np.ndarray, int, slice]):
        raise NotImplementedError

    @abstractmethod
    def shrink(self, new_lo: np.ndarray, new_up: np.ndarray):
        raise NotImplementedError

    @abstractmethod
    def contains(self, cand: np.ndarray, verbose: bool = False) -> bool:
        raise NotImplementedError

    def __contains__(self, cand):
        return self.contains(cand)

    @abstractmethod
    def sample_uniform(self, concrete_inf: float = 1e6) -> np.ndarray:
        raise NotImplementedError

    @abstractmethod
    def project_to(self, ele: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    @staticmethod
    def cat(spaces):
        raise NotImplementedError
