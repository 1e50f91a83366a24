"""
Reward functions (pyrado/tasks/reward_functions.py:176-297) as DESCRIPTORS: they hold Q, R (and min_rew) and compile
into the `sre_task` struct the kernels read.  Calling one evaluates it on the device through the C-ABI
(`sre_task_step_rew`); there is no host arithmetic on this path.
"""
from abc import ABC, abstractmethod

import numpy as np
import torch

from .. import _lib as L
from ..spaces.base import Space
from ..utils.exceptions import TypeErr, ValueErr


class RewFcn(ABC):
    rew_kind: int = None

    @abstractmethod
    def __call__(self, err_s: np.ndarray, err_a: np.ndarray, remaining_steps: int = None) -> float:
        raise NotImplementedError

    def reset(self, *args, **kwargs):
        """Called by Task.reset; only ScaledExpQuadrErrRewFcn has state (c_max)."""


class QuadrErrRewFcn(RewFcn):
    """rew = -(e_s^T Q e_s + e_a^T R e_a)   (reward_functions.py:176-222)"""

    rew_kind = L.REW_QUADR

    def __init__(self, Q, R):
        try:
            Q, R = np.array(Q, dtype=np.float64), np.array(R, dtype=np.float64)
        except Exception:
            raise TypeErr(given=Q, expected_type=[np.ndarray, list])
        Q, R = np.atleast_2d(Q), np.atleast_2d(R)
        if not (np.linalg.eigvals(Q).real >= 0).all():
            raise ValueErr(msg="The weight matrix Q must not have negative eigenvalues!")
        if not (np.linalg.eigvals(R).real >= 0).all():
            raise ValueErr(msg="The weight matrix R must not have negative eigenvalues!")
        if Q.shape[0] > L.MAX_DS or R.shape[0] > L.MAX_DA:
            raise ValueErr(msg=f"The engine supports at most {L.MAX_DS} state and {L.MAX_DA} action dimensions!")
        self.Q, self.R = Q, R
        self.c_max = 1.0
        self.min_rew = 1e-4

    def fill(self, task: L.Task):
        """Write this reward function into a C-ABI task struct."""
        ds, da = self.Q.shape[0], self.R.shape[0]
        for i in range(L.MAX_DS * L.MAX_DS):
            task.Q[i] = 0.0
        for i, v in enumerate(self.Q.reshape(-1)):
            task.Q[i] = v
        for i in range(L.MAX_DA * L.MAX_DA):
            task.R[i] = 0.0
        for i, v in enumerate(self.R.reshape(-1)):
            task.R[i] = v
        task.rew_kind = self.rew_kind
        task.min_rew = self.min_rew
        return ds, da

    def __call__(self, err_s: np.ndarray, err_a: np.ndarray, remaining_steps: int = None) -> float:
        assert isinstance(err_s, np.ndarray) and isinstance(err_a, np.ndarray)
        from .. import engine as E

        task = L.Task()
        ds, da = self.fill(task)  # state_des = 0, no modulation: cost(0 - (-e_s), e_a)
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else None
        if dev is None:
            raise L.EngineError("reward functions are evaluated on the CUDA device; none is available (no CPU fallback)")
        s = torch.as_tensor(-err_s.reshape(1, -1), dtype=torch.float64, device=dev).contiguous()
        a = torch.as_tensor(err_a.reshape(1, -1), dtype=torch.float64, device=dev).contiguous()
        return float(E.task_step_rew(task, ds, da, self.c_max, s, a)[0])


class ExpQuadrErrRewFcn(QuadrErrRewFcn):
    """rew = exp(-cost)   (reward_functions.py:224-244)"""

    rew_kind = L.REW_EXP


class ScaledExpQuadrErrRewFcn(QuadrErrRewFcn):
    """rew = exp(-c_max * cost), c_max = -ln(min_rew) / cost(worst case state, action)  (reward_functions.py:247-297).
    Inside the kernels c_max is recomputed per environment from its own (randomized) spaces at every episode start."""

    rew_kind = L.REW_SCALED_EXP

    def __init__(self, Q, R, state_space: Space, act_space: Space, min_rew: float = 1e-4):
        super().__init__(Q, R)
        self.reset(state_space, act_space, min_rew)

    def reset(self, state_space: Space, act_space: Space, min_rew=1e-4, **kwargs):
        if not isinstance(state_space, Space):
            raise TypeErr(given=state_space, expected_type=Space)
        if not isinstance(act_space, Space):
            raise TypeErr(given=act_space, expected_type=Space)
        self.min_rew = float(min_rew)
        s_max, a_max = state_space.bound_abs_up.reshape(-1), act_space.bound_abs_up.reshape(-1)
        max_cost = s_max.dot(self.Q.dot(s_max)) + a_max.dot(self.R.dot(a_max))
        self.c_max = float(-1.0 * np.log(self.min_rew) / max_cost)
