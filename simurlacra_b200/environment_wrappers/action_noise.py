"""
GaussianActNoiseWrapper (pyrado/environment_wrappers/action_noise.py:38-102): i.i.d. normal noise on every action on
its way into the environment, independent of a possible exploration strategy.  Fused into the kernels
(``cfg.env_act_noise``): the normals come from the Philox stream (seed, env, episode, stream 4, step); the wrapper may
sit outside or inside an ActNormWrapper (noise in normalised or in env units).  The recorded action stays the policy's.
"""
from typing import Optional, Union

import numpy as np
import torch

import simurlacra_b200 as pyrado

from .. import _lib as L
from .base import EnvWrapperAct


class GaussianActNoiseWrapper(EnvWrapperAct):
    """Environment wrapper which adds normally distributed i.i.d. noise to all actions."""

    _fusable = True

    def __init__(self, wrapped_env, noise_mean: Optional[Union[float, np.ndarray]] = None,
                 noise_std: Optional[Union[float, np.ndarray]] = None):
        super().__init__(wrapped_env)
        shape = self.act_space.shape
        self._mean = self._parse(noise_mean, shape)
        self._std = self._parse(noise_std, shape)

    def _parse(self, value, shape) -> np.ndarray:
        if value is None:
            return np.zeros(shape)
        arr = np.array(value, dtype=np.float64)
        if arr.shape == ():
            arr = np.full(shape, float(arr))  # the reference accepts a float in the signature; broadcast it
        if not arr.shape == shape:
            raise pyrado.ShapeErr(given=arr, expected_match=self.act_space)
        return arr

    def _ctor_kwargs(self) -> dict:
        return dict(noise_mean=self._mean, noise_std=self._std)

    def _fuse(self, fusion, below_alters_act, below_alters_obs) -> bool:
        from .action_normalization import ActNormWrapper

        if fusion.env_act_noise is not None:
            return False  # a second action-noise wrapper: run it on the data
        below = [w for w in self._below() if w._alters_act]
        if not below:
            pos = L.ENV_ACT_NOISE_INNER  # env units (inside an ActNormWrapper that was seen above, or none at all)
        elif len(below) == 1 and type(below[0]) is ActNormWrapper and not fusion.act_norm:
            pos = L.ENV_ACT_NOISE_OUTER  # normalised units, de-normalised afterwards
        else:
            return False
        fusion.env_act_noise = (pos, self._std, self._mean)
        return True

    def _process_act(self, act):
        """Un-fused path: host generator like the reference (:73-78)."""
        if isinstance(act, torch.Tensor):
            std, mean = torch.as_tensor(self._std).to(act), torch.as_tensor(self._mean).to(act)
            return act + (torch.randn_like(act) * std + mean)
        return act + (np.random.randn(*self.act_space.shape) * self._std + self._mean)

    def _set_wrapper_domain_param(self, domain_param: dict):
        domain_param["act_noise_mean"] = self._mean
        domain_param["act_noise_std"] = self._std

    def _get_wrapper_domain_param(self, domain_param: dict):
        for key, attr in (("act_noise_mean", "_mean"), ("act_noise_std", "_std")):
            if key in domain_param:
                arr = np.array(domain_param[key], dtype=np.float64)
                if not arr.shape == self.act_space.shape:
                    raise pyrado.ShapeErr(given=arr, expected_match=self.act_space)
                setattr(self, attr, arr)
