"""DiscreteSpace: a finite set of points, stored as an [N, D] table (behaviour of pyrado/spaces/discrete.py:39-175).

Inside the kernels only the three-point action set of BallOnBeamDiscSim exists (`SRE_F_DISC_ACT`, csrc/sre_core.cuh
`Lane::limit_act`); this class is the host-side descriptor the Python API hands out (`env.act_space`).  Semantics that
callers rely on and that the kernels reproduce: membership is `np.isclose` against any table row, and projecting a
non-member returns the row with the smallest absolute distance (first row on ties)."""
from typing import Optional, Sequence, Union

import numpy as np

from ..utils.exceptions import ShapeErr, TypeErr, ValueErr
from .base import Space


def _as_table(points) -> np.ndarray:
    """[N, D] table from an ndarray (kept in its dtype) or a list (integer table, like the reference's `np.intc`)."""
    if isinstance(points, list):
        table = np.asarray(points, dtype=np.intc)
    elif isinstance(points, np.ndarray):
        table = points
    else:
        raise TypeErr(given=points, expected_type=[np.ndarray, list])
    if table.ndim != 2:
        table = table.reshape(-1, 1)  # a flat list of scalars is N one-dimensional points
    return table


class DiscreteSpace(Space):
    """Finite space; `eles[i]` is the i-th point."""

    def __init__(self, eles: Union[np.ndarray, list], labels: Optional[Sequence[str]] = None):
        self.eles = _as_table(eles)
        self.bound_lo, self.bound_up = self.eles.min(axis=0), self.eles.max(axis=0)
        if labels is None:
            self._labels = np.full(self.shape, None, dtype=object)
        else:
            self._labels = np.array(labels, dtype=object)
            if self._labels.shape != self.shape:
                raise ShapeErr(given=self._labels, expected_match=self.bound_lo)

    def __str__(self):
        return f"DiscreteSpace(elements={self.eles.tolist()})"

    # ---- geometry -------------------------------------------------------------------------------------------------
    @property
    def shape(self) -> tuple:
        return (self.eles.shape[1],)

    @property
    def flat_dim(self) -> int:
        return self.eles.shape[1]

    @property
    def num_ele(self) -> int:
        return self.eles.shape[0]

    @property
    def bound_abs_up(self) -> np.ndarray:
        return np.abs(self.eles).max(axis=0)

    @property
    def labels(self) -> np.ndarray:
        return self._labels

    def _members(self) -> tuple:
        return self.eles, self._labels

    # ---- operations -------------------------------------------------------------------------------------------------
    def contains(self, cand: np.ndarray, verbose: bool = False) -> bool:
        cand = np.asarray(cand)
        if cand.shape != self.shape:
            raise ShapeErr(given=cand, expected_match=self.bound_lo)
        if np.isnan(cand).any():
            raise ValueErr(msg="At least one value is NaN!")
        hits = np.isclose(self.eles, cand.astype(self.eles.dtype))
        return bool(hits.any())

    def project_to(self, ele: np.ndarray) -> np.ndarray:
        if self.contains(ele):
            return ele
        nearest = int(np.argmin(np.abs(ele - self.eles)))  # flat index, like the reference; exact for D == 1
        return self.eles[nearest, :]

    def sample_uniform(self, concrete_inf: float = 1e6) -> np.ndarray:
        return self.eles[np.random.randint(self.num_ele), :].copy()

    def subspace(self, idcs):
        raise NotImplementedError("DiscreteSpace has no sub-spaces")

    def shrink(self, new_lo: np.ndarray, new_up: np.ndarray):
        raise NotImplementedError("DiscreteSpace cannot be shrunk")

    @staticmethod
    def cat(spaces):
        """Union of the point tables of several discrete spaces (`None` entries are skipped)."""
        tables = []
        for sp in spaces:
            if sp is None:
                continue
            if not isinstance(sp, DiscreteSpace):
                raise TypeErr(given=sp, expected_type=DiscreteSpace)
            tables.append(sp.eles)
        return DiscreteSpace(np.concatenate(tables, axis=0))
