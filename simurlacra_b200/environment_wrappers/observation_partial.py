"""
ObsPartialWrapper (pyrado/environment_wrappers/observation_partial.py:36-75): drop (or keep) selected observation
entries; the observation space shrinks accordingly.  One env (ndarray) or a batch ([num_envs, obs_dim] tensor).
"""
from typing import Optional

import numpy as np
import torch

import simurlacra_b200 as pyrado

from ..spaces.base import Space
from .base import EnvWrapperObs


class ObsPartialWrapper(EnvWrapperObs):
    """Environment wrapper which creates a partial observation by masking certain elements"""

    def __init__(self, wrapped_env, mask: Optional[list] = None, idcs: Optional[list] = None, keep_selected: bool = False):
        """
        :param mask: entries with 1 are dropped (inverted by `keep_selected=True`)
        :param idcs: indices (or labels of the observation space) to drop, ignored if `mask` is given
        :param keep_selected: keep the selected entries and drop the others
        """
        super().__init__(wrapped_env)
        self._ctor = dict(mask=mask, idcs=idcs, keep_selected=keep_selected)
        if mask is not None:
            mask = np.array(mask, dtype=bool)
            if not mask.shape == wrapped_env.obs_space.shape:
                raise pyrado.ShapeErr(given=mask, expected_match=wrapped_env.obs_space)
        else:
            assert idcs is not None, "Either mask or indices must be specified"
            mask = wrapped_env.obs_space.create_mask(idcs)
        self.keep_mask = mask if keep_selected else np.logical_not(mask)
        self._keep_idx = None

    def _fuse_rollout(self, fusion, below) -> bool:
        return True  # the persistent rollout kernel keeps this wrapper's state per lane (sampling/rollout.py::_compile_step_wrappers)

    def _ctor_kwargs(self) -> dict:
        return dict(self._ctor)

    def _process_obs(self, obs):
        if isinstance(obs, torch.Tensor):
            if self._keep_idx is None or self._keep_idx.device != obs.device:
                self._keep_idx = torch.as_tensor(np.flatnonzero(self.keep_mask), device=obs.device)
            return obs.index_select(-1, self._keep_idx)
        return obs[..., self.keep_mask]

    def _process_obs_space(self, space: Space) -> Space:
        return space.subspace(self.keep_mask)
