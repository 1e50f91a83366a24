"""QQubeSwingUpSim -- Quanser Qube Furuta pendulum (pyrado/environments/pysim/quanser_qube.py:43-188).
Dynamics, reward, done and observation run in the CUDA kernels (csrc/sre_envs.cuh, Env<SRE_QQ>)."""
import numpy as np

from ... import _lib as L
from ...spaces.box import BoxSpace
from ...tasks.desired_state import RadiallySymmDesStateTask
from ...tasks.reward_functions import ExpQuadrErrRewFcn
from .base import SimPyEnv

MAX_ACT_QQ = 4.5  # [V] environments/quanser/__init__.py:34


class QQubeSwingUpSim(SimPyEnv):
    name: str = "qq-su"
    kind = L.QQ
    _param_names = ("gravity_const", "motor_resistance", "motor_back_emf", "mass_rot_pole", "length_rot_pole",
                    "damping_rot_pole", "mass_pend_pole", "length_pend_pole", "damping_pend_pole", "voltage_thold_neg",
                    "voltage_thold_pos")

    @classmethod
    def get_nominal_domain_param(cls) -> dict:  # quanser_qube.py:54-68
        return dict(gravity_const=9.81, motor_resistance=8.4, motor_back_emf=0.042, mass_rot_pole=0.095,
                    length_rot_pole=0.085, damping_rot_pole=5e-6, mass_pend_pole=0.024, length_pend_pole=0.129,
                    damping_pend_pole=1e-6, voltage_thold_neg=0.0, voltage_thold_pos=0.0)

    def _create_spaces(self):  # :167-181
        max_state = np.array([115.0 / 180 * np.pi, 4 * np.pi, 20 * np.pi, 20 * np.pi])
        max_init_state = np.array([2.0, 1.0, 0.5, 0.5]) / 180 * np.pi
        max_obs = np.array([1.0, 1.0, 1.0, 1.0, 20 * np.pi, 20 * np.pi])
        self._state_space = BoxSpace(-max_state, max_state, labels=["theta", "alpha", "theta_dot", "alpha_dot"])
        self._obs_space = BoxSpace(-max_obs, max_obs, labels=["sin_theta", "cos_theta", "sin_alpha", "cos_alpha",
                                                              "theta_dot", "alpha_dot"])
        if not self._custom_init_space:
            self._init_space = BoxSpace(-max_init_state, max_init_state, labels=["theta", "alpha", "theta_dot", "alpha_dot"])
        self._act_space = BoxSpace(-MAX_ACT_QQ, MAX_ACT_QQ, shape=(1,), labels=["V"])

    def _create_task(self, task_args: dict):  # :183-188
        state_des = np.asarray(task_args.get("state_des", np.array([0.0, np.pi, 0.0, 0.0])), dtype=np.float64)
        Q = task_args.get("Q", np.diag([1.0, 1.0, 2e-2, 5e-3]))
        R = task_args.get("R", np.diag([4e-3]))
        return RadiallySymmDesStateTask(self.spec, state_des, ExpQuadrErrRewFcn(Q, R), idcs=[1])


class QQubeStabSim(QQubeSwingUpSim):
    """Stabilise the pendulum at the top (pyrado/environments/pysim/quanser_qube.py:191-229): the Qube's dynamics and
    state space with an init space around the upright pose (alpha in [175, 185] deg) and a stiffer reward."""

    name: str = "qq-st"
    _pass_init_box = True

    def _create_spaces(self):  # :203-217
        super()._create_spaces()
        if not self._custom_init_space:
            min_init_state = np.array([-5.0 / 180 * np.pi, 175.0 / 180 * np.pi, 0, 0])
            max_init_state = np.array([5.0 / 180 * np.pi, 185.0 / 180 * np.pi, 0, 0])
            self._init_space = BoxSpace(min_init_state, max_init_state, labels=["theta", "alpha", "theta_dot", "alpha_dot"])

    def _create_task(self, task_args: dict):  # :219-229
        state_des = np.asarray(task_args.get("state_des", np.array([0.0, np.pi, 0.0, 0.0])), dtype=np.float64)
        Q = task_args.get("Q", np.diag([3.0, 4.0, 2.0, 2.0]))
        R = task_args.get("R", np.diag([5e-2]))
        return RadiallySymmDesStateTask(self.spec, state_des, ExpQuadrErrRewFcn(Q, R), idcs=[1])
