"""PendulumSim -- under-actuated pendulum similar to the one from OpenAI Gym (pyrado/environments/pysim/pendulum.py:43-140).
Dynamics, reward, done and observation run in the CUDA kernels (csrc/sre_envs.cuh, Env<SRE_PEND>)."""
from typing import Optional

import numpy as np

import simurlacra_b200 as pyrado

from ... import _lib as L
from ...spaces.box import BoxSpace
from ...spaces.singular import SingularStateSpace
from ...tasks.desired_state import RadiallySymmDesStateTask
from ...tasks.reward_functions import ExpQuadrErrRewFcn
from .base import SimPyEnv


class PendulumSim(SimPyEnv):
    name: str = "pend"
    kind = L.PEND
    _param_names = ("gravity_const", "pole_mass", "pole_length", "pole_damping", "torque_thold")
    _pass_init_box = True  # the init space is the singular state given to the constructor

    def __init__(self, dt: float, max_steps: int = pyrado.inf, task_args: Optional[dict] = None,
                 init_state: Optional[np.ndarray] = None, **kwargs):
        """
        :param init_state: pole angle and angular velocity for the `SingularStateSpace` the episodes start from
        """
        self._init_state = np.zeros(2) if init_state is None else np.asarray(init_state, dtype=np.float64)  # [rad, rad/s]
        if self._init_state.size != 2:
            raise pyrado.ShapeErr(given=self._init_state, expected_match=(2,))
        super().__init__(dt, max_steps, task_args, **kwargs)

    @classmethod
    def get_nominal_domain_param(cls) -> dict:  # :96-104
        return dict(gravity_const=9.81, pole_mass=1.0, pole_length=1.0, pole_damping=0.05, torque_thold=3.5)

    def _create_spaces(self):  # :66-78 -- the action bound is the torque threshold, a domain parameter
        max_state = np.array([4 * np.pi, 4 * np.pi])
        max_obs = np.array([1.0, 1.0, np.inf])
        tau_max = self._domain_param["torque_thold"]
        self._state_space = BoxSpace(-max_state, max_state, labels=["theta", "theta_dot"])
        self._obs_space = BoxSpace(-max_obs, max_obs, labels=["sin_theta", "cos_theta", "theta_dot"])
        if not self._custom_init_space:
            self._init_space = SingularStateSpace(self._init_state, labels=["theta", "theta_dot"])
        self._act_space = BoxSpace(-np.array([tau_max]), np.array([tau_max]), labels=["tau"])

    def _create_task(self, task_args: dict):  # :80-88 (idcs=[1] as in the reference)
        state_des = np.asarray(task_args.get("state_des", np.array([np.pi, 0.0])), dtype=np.float64)
        Q = task_args.get("Q", np.diag([1e-0, 1e-3]))
        R = task_args.get("R", np.diag([1e-2]))
        return RadiallySymmDesStateTask(self.spec, state_des, ExpQuadrErrRewFcn(Q, R), idcs=[1])
