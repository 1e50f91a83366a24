"""
EnvWrapper / EnvWrapperAct / EnvWrapperObs (pyrado/environment_wrappers/base.py:45-385).

Delegation works like in the reference.  In addition every wrapper can *fuse* itself into the CUDA kernels of the
wrapped ``SimPyEnv``: calls travel down the chain through ``_reset_chain`` / ``_step_chain`` together with a
``Fusion`` record; a wrapper that the kernels implement (ActNormWrapper, GaussianObsNoiseWrapper) switches its flag on
instead of touching the data, everything else applies its ``_process_*`` hook to the arrays / tensors as in the
reference.  ``compile_chain`` gives samplers the (inner env, fusion, randomizer) triple for the persistent rollout
kernel, or ``None`` when some wrapper in the chain is not fusable (then they step through this generic path).
"""
from abc import abstractmethod
from typing import Iterable, Optional, Union

import numpy as np
import torch

import simurlacra_b200 as pyrado

from ..environments.base import Env
from ..environments.pysim.base import Fusion, SimPyEnv
from ..environments.sim_base import SimEnv
from ..spaces.base import Space
from ..tasks.base import Task
from ..utils.data_types import RenderMode


class EnvWrapper(Env):
    """Base for all environment wrappers. Delegates all environment methods to the wrapped environment."""

    _alters_act = False  # this wrapper changes the action on its way down
    _alters_obs = False  # this wrapper changes the observation on its way up
    _fusable = True      # transparent for the fused rollout kernel (pure delegation counts as fusable)

    def __init__(self, wrapped_env: Env):
        if not isinstance(wrapped_env, Env):
            raise pyrado.TypeErr(given=wrapped_env, expected_type=Env)
        self._wrapped_env = wrapped_env

    # ---- Serializable-like copy used by remove_env / insert_env_before (environment_wrappers/utils.py)
    def _ctor_kwargs(self) -> dict:
        return dict()

    def copy(self, **overrides) -> "EnvWrapper":
        kw = dict(wrapped_env=self._wrapped_env, **self._ctor_kwargs())
        kw.update(overrides)
        return type(self)(**kw)

    # ---- plain delegation (base.py:64-180)
    @property
    def name(self) -> str:
        return self._wrapped_env.name

    @property
    def wrapped_env(self) -> Env:
        return self._wrapped_env

    @property
    def num_envs(self) -> int:
        return getattr(self._wrapped_env, "num_envs", 1)

    @property
    def state_space(self) -> Space:
        return self._wrapped_env.state_space

    @property
    def obs_space(self) -> Space:
        return self._wrapped_env.obs_space

    @property
    def act_space(self) -> Space:
        return self._wrapped_env.act_space

    @property
    def init_space(self) -> Space:
        if isinstance(self._wrapped_env, (SimEnv, EnvWrapper)):
            return self._wrapped_env.init_space
        raise NotImplementedError

    @init_space.setter
    def init_space(self, space: Space):
        if isinstance(self._wrapped_env, (SimEnv, EnvWrapper)):
            self._wrapped_env.init_space = space
        else:
            raise NotImplementedError

    @property
    def dt(self):
        return self._wrapped_env.dt

    @dt.setter
    def dt(self, dt: float):
        self._wrapped_env.dt = dt

    @property
    def curr_step(self):
        return self._wrapped_env.curr_step

    @property
    def max_steps(self) -> Union[int, float]:
        return self._wrapped_env.max_steps

    @max_steps.setter
    def max_steps(self, num_steps: int):
        self._wrapped_env.max_steps = num_steps

    @property
    def state(self):
        """A copy of the wrapped environment's state (base.py:120-122)."""
        s = self._wrapped_env.state
        return s.clone() if isinstance(s, torch.Tensor) else s.copy()

    @state.setter
    def state(self, state):
        if not isinstance(state, (np.ndarray, torch.Tensor)):
            raise pyrado.TypeErr(given=state, expected_type=np.ndarray)
        cur = self._wrapped_env.state
        if cur is not None and tuple(state.shape) != tuple(cur.shape):
            raise pyrado.ShapeErr(given=state, expected_match=cur)
        self._wrapped_env.state = state

    def _create_task(self, task_args: dict) -> Task:
        return self._wrapped_env._create_task(task_args)

    @property
    def task(self) -> Task:
        return self._wrapped_env.task

    @property
    def domain_param(self) -> dict:
        if not isinstance(self._wrapped_env, (SimEnv, EnvWrapper)):
            raise pyrado.TypeErr(given=self._wrapped_env, expected_type=(SimEnv, EnvWrapper))
        param = self._wrapped_env.domain_param
        self._set_wrapper_domain_param(param)
        return param

    @domain_param.setter
    def domain_param(self, domain_param: dict):
        self._get_wrapper_domain_param(domain_param)
        self._wrapped_env.domain_param = domain_param

    def get_nominal_domain_param(self) -> dict:
        if not isinstance(self._wrapped_env, (SimEnv, EnvWrapper)):
            raise pyrado.TypeErr(given=self._wrapped_env, expected_type=(SimEnv, EnvWrapper))
        return self._wrapped_env.get_nominal_domain_param()

    @property
    def supported_domain_param(self) -> Iterable:
        return self._wrapped_env.supported_domain_param

    def forward(self, value):
        fcn = getattr(self._wrapped_env, "forward", None)
        return fcn(value) if callable(fcn) else value

    def inverse(self, value):
        fcn = getattr(self._wrapped_env, "inverse", None)
        return fcn(value) if callable(fcn) else value

    @property
    def randomizer(self):
        return getattr(self._wrapped_env, "randomizer", None)

    def _get_wrapper_domain_param(self, param: dict):
        """Called by the domain_param setter: load wrapper-specific entries. Nothing by default."""

    def _set_wrapper_domain_param(self, param: dict):
        """Called by the domain_param getter: store wrapper-specific entries. Nothing by default."""

    # ---- fusion protocol
    def _fuse(self, fusion: Fusion, below_alters_act: bool, below_alters_obs: bool) -> bool:
        """Record this wrapper in `fusion` if the kernels implement it; return False if it has to run on the arrays."""
        return True  # pure delegation needs nothing

    def _fuse_rollout(self, fusion: Fusion, below: list) -> bool:
        """The same question for the persistent rollout kernel, which also runs the stateful step wrappers
        (sampling/rollout.py::_compile_step_wrappers decides about their order); default: what `_fuse` says."""
        return self._fuse(fusion, any(b._alters_act for b in below), any(b._alters_obs for b in below))

    def _below(self):
        env = self._wrapped_env
        while isinstance(env, EnvWrapper):
            yield env
            env = env._wrapped_env

    def _try_fuse(self, fusion: Optional[Fusion]) -> bool:
        if fusion is None or not isinstance(inner_env_of(self), SimPyEnv):
            return False
        return self._fuse(fusion, any(w._alters_act for w in self._below()), any(w._alters_obs for w in self._below()))

    def _reset_chain(self, init_state, domain_param, fusion: Optional[Fusion]):
        if domain_param is not None:
            self._get_wrapper_domain_param(domain_param)
        self._try_fuse(fusion)
        return _call_reset(self._wrapped_env, init_state, domain_param, fusion)

    def _step_chain(self, act, fusion: Optional[Fusion]) -> tuple:
        self._try_fuse(fusion)
        return _call_step(self._wrapped_env, act, fusion)

    # ---- reset / step (base.py:220-247)
    def reset(self, init_state=None, domain_param: dict = None):
        return self._reset_chain(init_state, domain_param, Fusion())

    def step(self, act) -> tuple:
        return self._step_chain(act, Fusion())

    def step_diff_state(self, obs, act, act_diff, th_ddot=None):
        return self._wrapped_env.step_diff_state(obs, act, act_diff, th_ddot)

    def step_diff(self, obs, act, th_ddot=None):
        """Observation-based differentiable step, forwarded down the chain (environment_wrappers/base.py:245-256)."""
        return self._wrapped_env.step_diff(obs, act, th_ddot)

    def render(self, mode: RenderMode = None, render_step: int = 1):
        self._wrapped_env.render(mode, render_step)

    def close(self):
        return self._wrapped_env.close()


def inner_env_of(env):
    while isinstance(env, EnvWrapper):
        env = env._wrapped_env
    return env


def _call_reset(env, init_state, domain_param, fusion):
    if hasattr(env, "_reset_chain"):
        return env._reset_chain(init_state, domain_param, fusion)
    return env.reset(init_state=init_state, domain_param=domain_param)  # foreign Env (e.g. a mock)


def _call_step(env, act, fusion):
    if hasattr(env, "_step_chain"):
        return env._step_chain(act, fusion)
    return env.step(act)


class EnvWrapperAct(EnvWrapper):
    """Base class for wrappers modifying the action: override `_process_act` (and `_process_act_space`)."""

    _alters_act = True
    _fusable = False

    def _fuse(self, fusion, below_alters_act, below_alters_obs) -> bool:
        return False  # generic action wrappers run `_process_act` on the data

    @abstractmethod
    def _process_act(self, act):
        raise NotImplementedError

    def _process_act_tensor(self, act: torch.Tensor) -> torch.Tensor:
        return self._process_act(act)

    def _process_act_space(self, space: Space) -> Space:
        return space

    def _step_chain(self, act, fusion: Optional[Fusion]) -> tuple:
        if not self._try_fuse(fusion):
            act = self._process_act(act)
        return _call_step(self._wrapped_env, act, fusion)

    def step_diff(self, obs, act, th_ddot=None):
        """base.py:320-322: the action is processed, the observation passes through."""
        return self._wrapped_env.step_diff(obs, self._process_act_tensor(act), th_ddot)

    def step_diff_state(self, obs, act, act_diff, th_ddot=None):
        """base.py:324-327: both the action and the action offset are processed."""
        return self._wrapped_env.step_diff_state(obs, self._process_act_tensor(act), self._process_act_tensor(act_diff),
                                                 th_ddot)

    @property
    def act_space(self) -> Space:
        return self._process_act_space(self._wrapped_env.act_space)


class EnvWrapperObs(EnvWrapper):
    """Base class for wrappers modifying the observation: override `_process_obs` (and `_process_obs_space`)."""

    _alters_obs = True
    _fusable = False

    def _fuse(self, fusion, below_alters_act, below_alters_obs) -> bool:
        return False  # generic observation wrappers run `_process_obs` on the data

    @abstractmethod
    def _process_obs(self, obs):
        raise NotImplementedError

    def _process_obs_space(self, space: Space) -> Space:
        return space

    @property
    def obs_space(self) -> Space:
        return self._process_obs_space(self._wrapped_env.obs_space)

    def _reset_chain(self, init_state, domain_param, fusion: Optional[Fusion]):
        if domain_param is not None:
            self._get_wrapper_domain_param(domain_param)
        fused = self._try_fuse(fusion)
        obs = _call_reset(self._wrapped_env, init_state, domain_param, fusion)
        return obs if fused else self._process_obs(obs)

    def _step_chain(self, act, fusion: Optional[Fusion]) -> tuple:
        fused = self._try_fuse(fusion)
        obs, rew, done, info = _call_step(self._wrapped_env, act, fusion)
        return (obs if fused else self._process_obs(obs)), rew, done, info
