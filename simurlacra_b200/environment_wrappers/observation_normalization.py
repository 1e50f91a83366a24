"""
ObsNormWrapper (pyrado/environment_wrappers/observation_normalization.py:41-128): maps the observation to [-1, 1] with
the bounds of the wrapped observation space, or explicit per-label overrides for unbounded entries.  Fused into the
kernels (``cfg.obs_norm``), applied after a GaussianObsNoiseWrapper that sits below it.
"""
from typing import Mapping, Optional

import numpy as np
import torch

import simurlacra_b200 as pyrado

from ..spaces.box import BoxSpace
from .base import EnvWrapperObs


class ObsNormWrapper(EnvWrapperObs):
    """Environment wrapper which normalizes the observation space using the bounds from the environment or
    hard-coded bounds, such that all values are in range [-1, 1]"""

    _fusable = True

    def __init__(self, wrapped_env, explicit_lb: Optional[Mapping[str, float]] = None,
                 explicit_ub: Optional[Mapping[str, float]] = None):
        super().__init__(wrapped_env)
        self.explicit_lb, self.explicit_ub = explicit_lb, explicit_ub
        wos = self.wrapped_env.obs_space
        lb, ub = wos.bounds
        self.ov_lb = ObsNormWrapper.override_bounds(lb, self.explicit_lb, wos.labels)
        self.ov_ub = ObsNormWrapper.override_bounds(ub, self.explicit_ub, wos.labels)
        if any(self.ov_lb == -pyrado.inf):
            raise pyrado.ValueErr(msg=f"At least one element of the lower bounds is (negative) infinite:\n"
                                      f"(overwritten) bound: {self.ov_lb}\nnames: {wos.labels}")
        if any(self.ov_ub == pyrado.inf):
            raise pyrado.ValueErr(msg=f"At least one element of the upper bound is (positive) infinite:\n"
                                      f"(overwritten) bound: {self.ov_ub}\nnames: {wos.labels}")

    def _ctor_kwargs(self) -> dict:
        return dict(explicit_lb=self.explicit_lb, explicit_ub=self.explicit_ub)

    @staticmethod
    def override_bounds(bounds: np.ndarray, override: Optional[Mapping[str, float]], names: np.ndarray) -> np.ndarray:
        """A copy of `bounds` with the labelled entries replaced (:83-117); without overrides the bounds as they are."""
        if not override:
            return bounds
        bc = bounds.copy()
        for idx, name in np.ndenumerate(names):
            ov = override.get(name)
            if ov is not None:
                bc[idx] = ov
            elif np.isinf(bc[idx]):
                raise pyrado.ValueErr(msg=f"The entry {name} of a bound is infinite and not overwritten."
                                          f"Cannot apply normalization!")
        return bc

    def _fuse(self, fusion, below_alters_act, below_alters_obs) -> bool:
        from .observation_noise import GaussianObsNoiseWrapper

        if fusion.obs_norm is not None:
            return False
        if any(w._alters_obs and type(w) is not GaussianObsNoiseWrapper for w in self._below()):
            return False  # only the fused noise may sit between this wrapper and the simulation
        fusion.obs_norm = (self.ov_lb, self.ov_ub)
        return True

    def _process_obs(self, obs):
        if isinstance(obs, torch.Tensor):
            lb, ub = torch.as_tensor(self.ov_lb).to(obs), torch.as_tensor(self.ov_ub).to(obs)
            return (obs - lb) / (ub - lb) * 2 - 1
        return (obs - self.ov_lb) / (self.ov_ub - self.ov_lb) * 2 - 1

    def _process_obs_space(self, space: BoxSpace) -> BoxSpace:
        if not isinstance(space, BoxSpace):
            raise NotImplementedError("Only implemented ObsNormWrapper._process_obs_space() for BoxSpace!")
        return BoxSpace(-np.ones(space.shape), np.ones(space.shape), labels=space.labels)
