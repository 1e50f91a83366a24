"""QBallBalancerSim -- Quanser 2-DoF ball balancer (pyrado/environments/pysim/quanser_ball_balancer.py:50-444).
The plate angles (hidden state) are initialised at every reset by the 100-iteration float32 SGD inverse kinematics
(`QBallBalancerKin`, :340-444), which runs inside the reset kernel (csrc/sre_envs.cuh, Env<SRE_QBB>::plate_ang)."""
from typing import Optional

import numpy as np

import simurlacra_b200 as pyrado

from ... import _lib as L
from ...spaces.box import BoxSpace
from ...spaces.polar import Polar2DPosVelSpace
from ...tasks.desired_state import DesStateTask
from ...tasks.reward_functions import ScaledExpQuadrErrRewFcn
from .base import SimPyEnv

MAX_ACT_QBB = 3.0  # [V]


class QBallBalancerSim(SimPyEnv):
    name: str = "qbb"
    kind = L.QBB
    _param_names = ("gravity_const", "ball_mass", "ball_radius", "plate_length", "arm_radius", "gear_ratio",
                    "gear_efficiency", "load_inertia", "motor_inertia", "motor_back_emf", "motor_resistance",
                    "motor_efficiency", "combined_damping", "ball_damping", "voltage_thold_x_pos", "voltage_thold_x_neg",
                    "voltage_thold_y_pos", "voltage_thold_y_neg", "offset_th_x", "offset_th_y")

    def __init__(self, dt: float, max_steps: int = pyrado.inf, task_args: Optional[dict] = None,
                 simple_dynamics: bool = False, load_experimental_tholds: bool = True, **kwargs):
        super().__init__(dt, max_steps, task_args, **kwargs)
        self._simple_dynamics = simple_dynamics
        self._flags = L.F_SIMPLE_DYN if simple_dynamics else 0

    @classmethod
    def get_voltage_tholds(cls, load_experiments: bool = True) -> dict:
        """Hard-coded defaults (:135-138); the reference's optional measured-threshold files are not shipped."""
        return dict(voltage_thold_x_pos=0.28, voltage_thold_x_neg=-0.10, voltage_thold_y_pos=0.28, voltage_thold_y_neg=-0.074)

    @classmethod
    def get_nominal_domain_param(cls) -> dict:  # :169-202
        t = cls.get_voltage_tholds()
        return dict(gravity_const=9.81, ball_mass=0.003, ball_radius=0.019625, plate_length=0.275, arm_radius=0.0254,
                    gear_ratio=70.0, gear_efficiency=0.9, load_inertia=5.2822e-5, motor_inertia=4.6063e-7,
                    motor_back_emf=0.0077, motor_resistance=2.6, motor_efficiency=0.69, combined_damping=0.015,
                    ball_damping=0.05, voltage_thold_x_pos=t["voltage_thold_x_pos"],
                    voltage_thold_x_neg=t["voltage_thold_x_neg"], voltage_thold_y_pos=t["voltage_thold_y_pos"],
                    voltage_thold_y_neg=t["voltage_thold_y_neg"], offset_th_x=0.0, offset_th_y=0.0)

    def _create_spaces(self):  # :93-117
        l_plate = self._domain_param["plate_length"]
        max_state = np.array([np.pi / 4.0, np.pi / 4.0, l_plate / 2.0, l_plate / 2.0, 5 * np.pi, 5 * np.pi, 0.5, 0.5])
        min_init = np.array([0.75 * l_plate / 2, -np.pi, -0.05 * max_state[6], -0.05 * max_state[7]])
        max_init = np.array([0.8 * l_plate / 2, np.pi, 0.05 * max_state[6], 0.05 * max_state[7]])
        self._state_space = BoxSpace(-max_state, max_state, labels=["theta_x", "theta_y", "x", "y", "theta_x_dot",
                                                                    "theta_y_dot", "x_dot", "y_dot"])
        self._obs_space = self._state_space.copy()
        if not self._custom_init_space:
            self._init_space = Polar2DPosVelSpace(min_init, max_init, labels=["r", "phi", "x_dot", "y_dot"])
        self._act_space = BoxSpace(-MAX_ACT_QBB, MAX_ACT_QBB, shape=(2,), labels=["V_x", "V_y"])

    def _create_task(self, task_args: dict):  # :121-131
        state_des = np.asarray(task_args.get("state_des", np.zeros(8)), dtype=np.float64)
        Q = task_args.get("Q", np.diag([1e0, 1e0, 5e3, 5e3, 1e-2, 1e-2, 5e-1, 5e-1]))
        R = task_args.get("R", np.diag([1e-2, 1e-2]))
        return DesStateTask(self.spec, state_des,
                            ScaledExpQuadrErrRewFcn(Q, R, self.state_space, self.act_space, min_rew=1e-4))

    @property
    def plate_angs(self):
        h = self.hidden_state
        return None if h is None else (h[0].double().cpu().numpy() if self.num_envs == 1 else h)
