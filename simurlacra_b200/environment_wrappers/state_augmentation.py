"""
StateAugmentationWrapper (pyrado/environment_wrappers/state_augmentation.py:38-112): the observation is extended by
(selected) domain parameters of the inner simulation -- their current values, or the nominal ones if `fixed`.  For a
batch env every env appends its own parameter row.
"""
from typing import Optional, Sequence

import numpy as np
import torch

from ..spaces.box import BoxSpace
from .base import EnvWrapper, _call_reset, _call_step
from .utils import inner_env


class StateAugmentationWrapper(EnvWrapper):
    """Augments the observation of the wrapped environment by its physics configuration"""

    _alters_obs = True
    _fusable = False

    def __init__(self, wrapped_env, domain_param: Optional[Sequence[str]] = None, fixed: bool = False):
        """
        :param domain_param: names of the domain parameters to append, `None` selects all
        :param fixed: append the nominal values instead of the current ones
        """
        super().__init__(wrapped_env)
        sim = inner_env(wrapped_env)
        self._params = list(domain_param) if domain_param is not None else list(sim.domain_param.keys())
        nominal = sim.get_nominal_domain_param()
        self._nominal = np.array([nominal[k] for k in self._params], dtype=np.float64)
        self.fixed = fixed

    def _fuse_rollout(self, fusion, below) -> bool:
        return True  # the persistent rollout kernel keeps this wrapper's state per lane (sampling/rollout.py::_compile_step_wrappers)

    def _ctor_kwargs(self) -> dict:
        return dict(domain_param=self._params, fixed=self.fixed)

    def _fuse(self, fusion, below_alters_act, below_alters_obs) -> bool:
        return False

    def _params_as_tensor(self):
        if self.fixed:
            return self._nominal
        dp = inner_env(self._wrapped_env).domain_param
        return np.array([dp[k] for k in self._params], dtype=np.float64)  # [k] or [k, num_envs]

    def _augment(self, obs):
        prm = self._params_as_tensor()
        if isinstance(obs, torch.Tensor):
            p = torch.as_tensor(prm, dtype=obs.dtype, device=obs.device)
            p = p.t() if p.dim() == 2 else p.expand(*obs.shape[:-1], p.shape[0])
            return torch.cat((obs, p), dim=-1)
        return np.concatenate((obs, prm))

    @property
    def obs_space(self):
        lo, up = np.minimum(0.5 * self._nominal, 1.5 * self._nominal), np.maximum(0.5 * self._nominal, 1.5 * self._nominal)
        return BoxSpace.cat((self._wrapped_env.obs_space, BoxSpace(lo, up, labels=self._params)))

    def _step_chain(self, act, fusion) -> tuple:
        obs, rew, done, info = _call_step(self._wrapped_env, act, None)
        return self._augment(obs), rew, done, info

    def _reset_chain(self, init_state, domain_param, fusion):
        if domain_param is not None:
            self._get_wrapper_domain_param(domain_param)
        return self._augment(_call_reset(self._wrapped_env, init_state, domain_param, None))

    @property
    def mask(self) -> np.ndarray:
        return np.concatenate((np.zeros(self._wrapped_env.obs_space.flat_dim), np.ones(len(self._params))))

    @property
    def offset(self) -> int:
        return self._wrapped_env.obs_space.flat_dim

    @property
    def param(self) -> dict:
        return inner_env(self._wrapped_env).domain_param

    @param.setter
    def param(self, params):
        inner_env(self._wrapped_env).domain_param = {k: float(v) for k, v in zip(self._params, params)}

    def set_adv(self, params):
        """Nominal value + offset for every augmented parameter (the reference's :103-105 indexes its nominal ARRAY by
        name, which raises; the intended update is implemented)."""
        inner_env(self._wrapped_env).domain_param = {k: float(n + v) for k, n, v in zip(self._params, self._nominal, params)}

    @property
    def nominal(self) -> np.ndarray:
        return self._nominal
