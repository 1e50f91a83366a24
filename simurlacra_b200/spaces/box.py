"""BoxSpace (pyrado/spaces/box.py:39-226): inclusive bounds, NaN candidates raise ValueErr (:142), project_to clips."""
from typing import Sequence, Union

import numpy as np

from ..utils.exceptions import ShapeErr, TypeErr, ValueErr
from .base import Space


class BoxSpace(Space):
    def __init__(self, bound_lo, bound_up, shape: Union[tuple, int] = None, labels: Sequence[str] = None):
        lo, up = np.asarray(bound_lo, dtype=np.float64), np.asarray(bound_up, dtype=np.float64)
        if shape is not None:  # scalar bounds broadcast to the given shape
            lo, up = np.full(shape, lo, dtype=np.float64), np.full(shape, up, dtype=np.float64)
        lo, up = np.atleast_1d(lo).copy(), np.atleast_1d(up).copy()
        if lo.shape != up.shape:
            raise ShapeErr(given=up, expected_match=lo)
        if np.any(lo > up):
            raise ValueErr(msg="All lower bounds must be smaller or equal to the upper bounds!")
        self.bound_lo, self.bound_up = lo, up
        self._labels = None
        if labels is not None:
            self._labels = np.array(labels, dtype=object)
            if self._labels.shape != lo.shape:
                raise ShapeErr(given=self._labels, expected_match=lo)

    def _members(self):
        return self.bound_lo, self.bound_up, np.array([] if self._labels is None else self._labels, dtype=object)

    @property
    def shape(self) -> tuple:
        return self.bound_lo.shape

    @property
    def labels(self):
        return self._labels

    def subspace(self, idcs):
        if isinstance(idcs, (list, tuple)):
            idcs = np.asarray(idcs)
        lo, up = np.atleast_1d(self.bound_lo[idcs]), np.atleast_1d(self.bound_up[idcs])
        lab = None if self._labels is None else np.atleast_1d(self._labels[idcs])
        return BoxSpace(lo, up, labels=lab)

    def shrink(self, new_lo: np.ndarray, new_up: np.ndarray):
        new_lo, new_up = np.asarray(new_lo, dtype=np.float64), np.asarray(new_up, dtype=np.float64)
        if new_lo.shape != self.shape or new_up.shape != self.shape:
            raise ShapeErr(given=new_lo, expected_match=self.bound_lo)
        if np.any(new_lo < self.bound_lo) or np.any(new_up > self.bound_up):
            raise ValueErr(msg="The new bounds must lie inside the old ones!")
        return BoxSpace(new_lo, new_up, labels=self._labels)

    def contains(self, cand: np.ndarray, verbose: bool = False) -> bool:
        cand = np.asarray(cand)
        if cand.shape != self.shape:
            raise ShapeErr(given=cand, expected_match=self)
        if np.isnan(cand).any():
            raise ValueErr(msg=f"At least one value is NaN!\n{cand}")
        inside = bool(np.all(cand >= self.bound_lo) and np.all(cand <= self.bound_up))
        if not inside and verbose:
            bad = np.where((cand < self.bound_lo) | (cand > self.bound_up))[0]
            print(f"Violated bounds at indices {bad.tolist()}: {cand[bad]}")
        return inside

    def sample_uniform(self, concrete_inf: float = 1e6) -> np.ndarray:
        lo = np.where(np.isinf(self.bound_lo), -concrete_inf, self.bound_lo)
        up = np.where(np.isinf(self.bound_up), concrete_inf, self.bound_up)
        return np.random.uniform(lo, up)

    def project_to(self, ele: np.ndarray) -> np.ndarray:
        if hasattr(ele, "clamp") and hasattr(ele, "device"):  # batch of elements as a torch tensor [N, *shape]
            import torch

            lo = torch.as_tensor(self.bound_lo, dtype=ele.dtype, device=ele.device)
            up = torch.as_tensor(self.bound_up, dtype=ele.dtype, device=ele.device)
            return torch.maximum(torch.minimum(ele, up), lo)
        if not isinstance(ele, np.ndarray):
            raise TypeErr(given=ele, expected_type=np.ndarray)
        return ele if self.contains(ele) else np.clip(ele, self.bound_lo, self.bound_up)

    @staticmethod
    def cat(spaces):
        spaces = list(spaces)
        if not all(isinstance(s, BoxSpace) for s in spaces):
            raise TypeErr(given=spaces, expected_type=BoxSpace)
        lo = np.concatenate([s.bound_lo.reshape(-1) for s in spaces])
        up = np.concatenate([s.bound_up.reshape(-1) for s in spaces])
        labs = None
        if all(s.labels is not None for s in spaces):
            labs = np.concatenate([s.labels.reshape(-1) for s in spaces])
        return BoxSpace(lo, up, labels=labs)


class InfBoxSpace(BoxSpace):
    """Box with infinite bounds in every dimension (box.py:210-226)."""

    def __init__(self, shape: Union[tuple, int], labels: Sequence[str] = None):
        super().__init__(-np.inf, np.inf, shape=shape, labels=labels)
