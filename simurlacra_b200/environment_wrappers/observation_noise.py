"""
GaussianObsNoiseWrapper (pyrado/environment_wrappers/observation_noise.py:39-97): i.i.d. normal noise on every
observation, at reset and at every step.  Fused into the kernels (``cfg.obs_noise``): the normals come from the Philox
stream keyed (seed, env, episode, step), so a rollout is reproducible for any batch split.
"""
from typing import Optional, Union

import numpy as np
import torch

import simurlacra_b200 as pyrado

from .base import EnvWrapperObs


class GaussianObsNoiseWrapper(EnvWrapperObs):
    """Environment wrapper which adds normally distributed i.i.d. noise to all observations."""

    _fusable = True

    def __init__(self, wrapped_env, noise_std: Union[list, np.ndarray], noise_mean: Optional[Union[list, np.ndarray]] = None):
        super().__init__(wrapped_env)
        self._std = np.array(noise_std, dtype=np.float64)
        if not self._std.shape == self.obs_space.shape:
            raise pyrado.ShapeErr(given=self._std, expected_match=self.obs_space)
        if noise_mean is not None:
            self._mean = np.array(noise_mean, dtype=np.float64)
            if not self._mean.shape == self.obs_space.shape:
                raise pyrado.ShapeErr(given=self._mean, expected_match=self.obs_space)
        else:
            self._mean = np.zeros(self.obs_space.shape)

    def _ctor_kwargs(self) -> dict:
        return dict(noise_std=self._std, noise_mean=self._mean)

    def _fuse(self, fusion, below_alters_act, below_alters_obs) -> bool:
        if below_alters_obs or fusion.noise_std is not None:
            return False  # something reshapes the observation below, or a second noise wrapper: run on the data
        fusion.noise_std, fusion.noise_mean = self._std, self._mean
        return True

    def _process_obs(self, obs):
        """Un-fused path (a non-fusable observation wrapper sits below): host generator like the reference (:68-73)."""
        if isinstance(obs, torch.Tensor):
            std, mean = torch.as_tensor(self._std).to(obs), torch.as_tensor(self._mean).to(obs)
            return obs + torch.randn_like(obs) * std + mean
        return obs + np.random.randn(*self.obs_space.shape) * self._std + self._mean

    def _set_wrapper_domain_param(self, domain_param: dict):
        domain_param["obs_noise_mean"] = self._mean
        domain_param["obs_noise_std"] = self._std

    def _get_wrapper_domain_param(self, domain_param: dict):
        if "obs_noise_mean" in domain_param:
            self._mean = np.array(domain_param["obs_noise_mean"], dtype=np.float64)
            assert self._mean.shape == self.obs_space.shape
        if "obs_noise_std" in domain_param:
            self._std = np.array(domain_param["obs_noise_std"], dtype=np.float64)
            assert self._std.shape == self.obs_space.shape
