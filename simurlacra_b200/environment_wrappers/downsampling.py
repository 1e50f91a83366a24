"""
DownsamplingWrapper (pyrado/environment_wrappers/downsampling.py:38-162): every `factor`-th action is held for
`factor` steps; the observation handed to the policy is a filter (default: the mean) over the last `factor`
observations.  `downsampling` is the wrapper's domain parameter.  One env (ndarray) or a batch (device tensors).
"""
from collections import deque
from typing import Callable, Optional

import numpy as np
import torch

from .base import EnvWrapperAct, EnvWrapperObs, _call_reset, _call_step


class DownsamplingWrapper(EnvWrapperAct, EnvWrapperObs):
    """Environment wrapper which downsamples the actions coming from the rollout loop."""

    def __init__(self, wrapped_env, factor: int, obs_filter_fcn: Optional[Callable] = None,
                 init_obs: Optional[np.ndarray] = None):
        """
        :param factor: number of time steps for which every action is repeated
        :param obs_filter_fcn: operates along dimension 0 of the stacked buffer; `None` = mean (the reference's default
                               `functools.partial(np.mean, axis=0)`)
        :param init_obs: initial observation to seed the buffer with
        """
        super().__init__(wrapped_env)
        self._factor = factor
        self._cnt = 0
        self._act_last = None
        self._obs_buffer = deque([] if init_obs is None else [np.array(init_obs, copy=True)], maxlen=factor)
        self._obs_filter_fcn = obs_filter_fcn

    def _fuse_rollout(self, fusion, below) -> bool:
        return True  # the persistent rollout kernel keeps this wrapper's state per lane (sampling/rollout.py::_compile_step_wrappers)

    def _ctor_kwargs(self) -> dict:
        return dict(factor=self._factor, obs_filter_fcn=self._obs_filter_fcn)

    @property
    def factor(self) -> int:
        return self._factor

    @factor.setter
    def factor(self, factor: int):
        assert isinstance(factor, int) and factor >= 1
        self._factor = factor
        self._cnt = 0

    def _set_wrapper_domain_param(self, domain_param: dict):
        domain_param["downsampling"] = int(self._factor)

    def _get_wrapper_domain_param(self, domain_param: dict):
        self._factor = int(domain_param.get("downsampling", self._factor))

    def _process_act(self, act):
        if self._factor >= 1:
            if self._cnt % self._factor == 0:
                self._act_last = act.clone() if isinstance(act, torch.Tensor) else np.array(act, copy=True)
            else:
                act = self._act_last.clone() if isinstance(self._act_last, torch.Tensor) else self._act_last.copy()
            self._cnt += 1
        return act

    def _process_obs(self, obs):
        self._obs_buffer.append(obs)
        if isinstance(obs, torch.Tensor):
            stack = torch.stack(list(self._obs_buffer))
            return stack.mean(dim=0) if self._obs_filter_fcn is None else self._obs_filter_fcn(stack)
        return np.mean(self._obs_buffer, axis=0) if self._obs_filter_fcn is None else self._obs_filter_fcn(self._obs_buffer)

    def _reset_chain(self, init_state, domain_param, fusion):
        if domain_param is not None:
            self._get_wrapper_domain_param(domain_param)
        self._act_last, self._cnt = None, 0
        obs = _call_reset(self._wrapped_env, init_state, domain_param, None)
        # like the reference (:118-131 over base.py:368-373): the initial observation passes through the filter with
        # the buffer as the previous episode left it, and the result seeds the new buffer
        obs = self._process_obs(obs)
        self._obs_buffer = deque([obs.clone() if isinstance(obs, torch.Tensor) else obs.copy()], maxlen=self._factor)
        return obs

    def _step_chain(self, act, fusion) -> tuple:
        obs, rew, done, info = _call_step(self._wrapped_env, self._process_act(act), None)
        return self._process_obs(obs), rew, done, info
