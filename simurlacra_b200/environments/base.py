"""Env interface (pyrado/environments/base.py:43-236)."""
from abc import ABC, abstractmethod
from typing import Optional, Union

import numpy as np

import simurlacra_b200 as pyrado

from ..spaces.base import Space
from ..tasks.base import Task
from ..utils.data_types import EnvSpec, RenderMode


class Env(ABC):
    name: str = None  # unique identifier

    def __init__(self, dt: Union[int, float], max_steps: Union[int, float] = pyrado.inf):
        """Same argument contract as the reference (environments/base.py:55-60): a non-number `dt` is a TypeErr, a
        negative `dt` or `max_steps < 1` a ValueErr."""
        self._check_number(dt)
        if dt < 0:
            raise pyrado.ValueErr(given=dt, ge_constraint="0")
        if max_steps < 1:
            raise pyrado.ValueErr(given=max_steps, ge_constraint="1")
        self._dt, self._max_steps = float(dt), max_steps
        self._curr_step, self._curr_rew = 0, -pyrado.inf

    @staticmethod
    def _check_number(x):
        if not isinstance(x, (int, float)):
            raise pyrado.TypeErr(given=x, expected_type=(int, float))

    @property
    @abstractmethod
    def state_space(self) -> Space:
        raise NotImplementedError

    @property
    @abstractmethod
    def obs_space(self) -> Space:
        raise NotImplementedError

    @property
    @abstractmethod
    def act_space(self) -> Space:
        raise NotImplementedError

    @property
    def spec(self) -> EnvSpec:
        return EnvSpec(self.obs_space, self.act_space, self.state_space)

    @property
    def dt(self) -> float:
        return self._dt

    @dt.setter
    def dt(self, dt: Union[int, float]):
        self._check_number(dt)
        if dt <= 0:
            raise pyrado.ValueErr(given=dt, g_constraint="0")
        self._dt = float(dt)

    @property
    def max_steps(self) -> Union[int, float]:
        return self._max_steps

    @max_steps.setter
    def max_steps(self, num_steps: Union[int, float]):
        unbounded = num_steps == pyrado.inf
        if not (unbounded or isinstance(num_steps, int)):
            raise pyrado.TypeErr(msg=f"max_steps takes an int or pyrado.inf, got {num_steps!r}")
        if num_steps <= 0:
            raise pyrado.ValueErr(given=num_steps, g_constraint="0")
        self._max_steps = num_steps

    @property
    def curr_step(self):
        return self._curr_step

    @property
    @abstractmethod
    def task(self) -> Task:
        raise NotImplementedError

    @abstractmethod
    def reset(self, init_state: Optional[np.ndarray] = None, domain_param: Optional[dict] = None):
        raise NotImplementedError

    @abstractmethod
    def step(self, act) -> tuple:
        raise NotImplementedError

    def limit_act(self, act: np.ndarray) -> np.ndarray:
        return self.act_space.project_to(act)

    def render(self, mode: RenderMode = None, render_step: int = 1):
        """Rendering (Panda3D in the reference) is unavailable on this box: no graphics stack."""

    def close(self):
        pass
