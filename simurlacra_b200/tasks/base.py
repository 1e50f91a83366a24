"""Task interface (pyrado/tasks/base.py:40-290), reduced to what the pysim path uses."""
from abc import ABC, abstractmethod

import numpy as np

from ..utils.data_types import EnvSpec


class Task(ABC):
    @property
    @abstractmethod
    def env_spec(self) -> EnvSpec:
        raise NotImplementedError

    @abstractmethod
    def reset(self, env_spec: EnvSpec, **kwargs):
        raise NotImplementedError

    @abstractmethod
    def step_rew(self, state: np.ndarray, act: np.ndarray, remaining_steps: int) -> float:
        raise NotImplementedError

    def final_rew(self, state: np.ndarray, remaining_steps: int) -> float:
        """tasks/base.py:122-131: zero unless wrapped in a FinalRewTask"""
        return 0.0

    @abstractmethod
    def has_succeeded(self, state: np.ndarray) -> bool:
        raise NotImplementedError

    def has_failed(self, state: np.ndarray) -> bool:
        """tasks/base.py:159-167: failed <=> the state left the state space"""
        return not self.env_spec.state_space.contains(state)

    def is_done(self, state: np.ndarray) -> bool:
        """tasks/base.py:169-172"""
        return self.has_succeeded(state) or self.has_failed(state)

    @abstractmethod
    def fill(self, task_struct) -> None:
        """Compile this task into the C-ABI `sre_task` struct."""
        raise NotImplementedError


class TaskWrapper(Task):
    """tasks/base.py:215-290"""

    def __init__(self, wrapped_task: Task):
        if not isinstance(wrapped_task, Task):
            from ..utils.exceptions import TypeErr

            raise TypeErr(given=wrapped_task, expected_type=Task)
        self._wrapped_task = wrapped_task

    @property
    def env_spec(self) -> EnvSpec:
        return self._wrapped_task.env_spec

    @property
    def wrapped_task(self) -> Task:
        return self._wrapped_task

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self._wrapped_task, name)

    def reset(self, **kwargs):
        self._wrapped_task.reset(**kwargs)

    def step_rew(self, state, act, remaining_steps=None) -> float:
        return self._wrapped_task.step_rew(state, act, remaining_steps)

    def has_succeeded(self, state) -> bool:
        return self._wrapped_task.has_succeeded(state)

    def fill(self, task_struct) -> None:
        self._wrapped_task.fill(task_struct)
