"""DesStateTask (pyrado/tasks/desired_state.py:40-114) and RadiallySymmDesStateTask (:116-155) as descriptors with
device evaluation: `step_rew` runs `sre_task_step_rew` through the C-ABI."""
from typing import Callable, Sequence

import numpy as np
import torch

from .. import _lib as L
from ..utils.data_types import EnvSpec
from ..utils.exceptions import ShapeErr, TypeErr
from .base import Task
from .reward_functions import RewFcn
from .utils import never_succeeded


class DesStateTask(Task):
    def __init__(self, env_spec: EnvSpec, state_des: np.ndarray, rew_fcn: RewFcn, success_fcn: Callable = None):
        if not isinstance(env_spec, EnvSpec):
            raise TypeErr(given=env_spec, expected_type=EnvSpec)
        if not isinstance(state_des, np.ndarray):
            raise TypeErr(given=state_des, expected_type=np.ndarray)
        if not isinstance(rew_fcn, RewFcn):
            raise TypeErr(given=rew_fcn, expected_type=RewFcn)
        self._env_spec = env_spec
        self._state_des = state_des
        self._rew_fcn = rew_fcn
        self.success_fcn = success_fcn if success_fcn is not None else never_succeeded

    @property
    def env_spec(self) -> EnvSpec:
        return self._env_spec

    @property
    def state_des(self) -> np.ndarray:
        return self._state_des

    @state_des.setter
    def state_des(self, state_des: np.ndarray):
        if not isinstance(state_des, np.ndarray):
            raise TypeErr(given=state_des, expected_type=np.ndarray)
        if not state_des.shape == self.state_des.shape:
            raise ShapeErr(given=state_des, expected_match=self.state_des)
        self._state_des = state_des

    @property
    def rew_fcn(self) -> RewFcn:
        return self._rew_fcn

    @rew_fcn.setter
    def rew_fcn(self, rew_fcn: RewFcn):
        if not isinstance(rew_fcn, RewFcn):
            raise TypeErr(given=rew_fcn, expected_type=RewFcn)
        self._rew_fcn = rew_fcn

    def reset(self, env_spec: EnvSpec, state_des: np.ndarray = None, **kwargs):
        self._env_spec = env_spec  # spaces may have changed with the domain parameters
        if state_des is not None:
            self._state_des = state_des
        self._rew_fcn.reset(state_space=env_spec.state_space, act_space=env_spec.act_space, **kwargs)

    # ---- compilation into the C-ABI struct ----
    def fill(self, task: L.Task) -> None:
        for i in range(L.MAX_DS):
            task.state_des[i] = 0.0
            task.radial_mod[i] = 2 * np.pi
        for i, v in enumerate(np.asarray(self._state_des, dtype=np.float64).reshape(-1)):
            task.state_des[i] = v
        task.radial_mask = 0
        task.final_flags = 0
        task.final_factor = 0.0
        self._rew_fcn.fill(task)

    def step_rew(self, state: np.ndarray, act: np.ndarray, remaining_steps: int = None) -> float:
        from .. import engine as E

        if not torch.cuda.is_available():
            raise L.EngineError("tasks are evaluated on the CUDA device; none is available (no CPU fallback)")
        task = L.Task()
        self.fill(task)
        dev = torch.device("cuda", torch.cuda.current_device())
        s = torch.as_tensor(np.asarray(state, dtype=np.float64).reshape(1, -1), device=dev).contiguous()
        a = torch.as_tensor(np.asarray(act, dtype=np.float64).reshape(1, -1), device=dev).contiguous()
        return float(E.task_step_rew(task, s.shape[1], a.shape[1], getattr(self._rew_fcn, "c_max", 1.0), s, a)[0])

    def has_succeeded(self, state: np.ndarray) -> bool:
        return self.success_fcn(self.state_des - state)


class RadiallySymmDesStateTask(DesStateTask):
    """A subset of the state error is taken modulo `modulation` (fmod), then EVERY element above pi / below -pi is
    folded back (desired_state.py:146-155) -- the kernels reproduce exactly this."""

    def __init__(self, env_spec: EnvSpec, state_des: np.ndarray, rew_fcn: RewFcn, idcs: Sequence[int],
                 modulation=2 * np.pi, success_fcn: Callable = None):
        super().__init__(env_spec, state_des, rew_fcn, success_fcn)
        self.idcs = list(idcs)
        self.mod = modulation * np.ones(len(self.idcs))

    def fill(self, task: L.Task) -> None:
        super().fill(task)
        task.radial_mask = sum(1 << int(i) for i in self.idcs)
        for i, m in zip(self.idcs, self.mod):
            task.radial_mod[int(i)] = float(m)
