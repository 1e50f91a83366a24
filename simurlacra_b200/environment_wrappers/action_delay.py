"""
ActDelayWrapper (pyrado/environment_wrappers/action_delay.py:37-112): the action that reaches the environment is the
one commanded `delay` steps earlier; the queue starts with `delay` zero actions at every reset.  The delay is a domain
parameter (`act_delay`) and can be randomized.  Works on one env (ndarray) and on a batch ([num_envs, act_dim] device
tensor, one queue entry per step for the whole batch) between step-kernel launches.
"""
import numpy as np
import torch

import simurlacra_b200 as pyrado

from .base import EnvWrapperAct


class ActDelayWrapper(EnvWrapperAct):
    """Environment wrapper which delays actions by a fixed number of time steps."""

    def __init__(self, wrapped_env, delay: int = 0):
        super().__init__(wrapped_env)
        self._delay = delay
        self._act_queue = []

    def _fuse_rollout(self, fusion, below) -> bool:
        return True  # the persistent rollout kernel keeps this wrapper's state per lane (sampling/rollout.py::_compile_step_wrappers)

    def _ctor_kwargs(self) -> dict:
        return dict(delay=self._delay)

    @property
    def delay(self) -> int:
        """The (rounded) action delay measured in time steps."""
        if isinstance(self._delay, np.ndarray):
            return int(np.round(self._delay).reshape(-1)[0]) if self._delay.size == 1 else np.round(self._delay)
        return round(self._delay)

    @delay.setter
    def delay(self, delay: int):
        if not delay >= 0:
            raise pyrado.ValueErr(given=delay, ge_constraint="0")
        self._delay = round(delay)

    def _set_wrapper_domain_param(self, domain_param: dict):
        domain_param["act_delay"] = self._delay

    def _get_wrapper_domain_param(self, domain_param: dict):
        self._delay = domain_param.get("act_delay", self._delay)

    def _reset_chain(self, init_state, domain_param, fusion):
        obs = super()._reset_chain(init_state, domain_param, fusion)
        self._act_queue = [None] * int(self.delay)  # zero actions, materialised with the first action's type and shape
        return obs

    def _process_act(self, act):
        if int(self.delay) != 0:
            self._act_queue.append(act.clone() if isinstance(act, torch.Tensor) else np.array(act, copy=True))
            first = self._act_queue.pop(0)
            act = (torch.zeros_like(act) if isinstance(act, torch.Tensor) else np.zeros_like(act)) if first is None else first
        return act
