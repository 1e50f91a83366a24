"""FinalRewMode / FinalRewTask (pyrado/tasks/final_reward.py:41-275): a reward or cost added once when the episode
ends.  Compiled into `sre_task.final_flags` / `final_factor`; the kernels add it in the step that sets `done`."""
from typing import NamedTuple

import numpy as np

from .. import _lib as L
from ..utils.exceptions import TypeErr
from .base import Task, TaskWrapper


class FinalRewMode(NamedTuple):
    state_dependent: bool = False
    time_dependent: bool = False
    always_positive: bool = False
    always_negative: bool = False
    user_input: bool = False


class FinalRewTask(TaskWrapper):
    def __init__(self, wrapped_task: Task, mode: FinalRewMode, factor: float = 1e3):
        super().__init__(wrapped_task)
        if not isinstance(mode, FinalRewMode):
            raise TypeErr(given=mode, expected_type=FinalRewMode)
        if mode.user_input:
            raise NotImplementedError("FinalRewMode.user_input asks for console input per episode; not available in a kernel")
        self.mode = mode
        self.factor = abs(factor)
        self._yielded_final_rew = False

    @property
    def yielded_final_rew(self) -> bool:
        return self._yielded_final_rew

    def reset(self, **kwargs):
        super().reset(**kwargs)
        self._yielded_final_rew = False

    def fill(self, task: L.Task) -> None:
        self._wrapped_task.fill(task)
        m = self.mode
        task.final_flags = (L.FR_ENABLED | (L.FR_STATE_DEP if m.state_dependent else 0) | (L.FR_TIME_DEP if m.time_dependent else 0)
                            | (L.FR_ALWAYS_POS if m.always_positive else 0) | (L.FR_ALWAYS_NEG if m.always_negative else 0))
        task.final_factor = float(self.factor)

    def compute_final_rew(self, state: np.ndarray, remaining_steps: int) -> float:
        """Host view of the rule the kernels implement (success_fcn == never_succeeded in the pysim envs, so only the
        failure branches can fire, final_reward.py:111-275)."""
        if self._yielded_final_rew:
            return 0.0
        self._yielded_final_rew = True
        if self.has_succeeded(state) or not self.has_failed(state) or self.mode.always_positive:
            return 0.0
        scale = float(remaining_steps) if self.mode.time_dependent else self.factor
        if self.mode.state_dependent:
            act = np.zeros(self.env_spec.act_space.shape)
            return -1.0 * scale * abs(self.step_rew(state, act, remaining_steps))
        return -1.0 * scale

    def final_rew(self, state: np.ndarray, remaining_steps: int) -> float:
        return self.compute_final_rew(state, remaining_steps)
