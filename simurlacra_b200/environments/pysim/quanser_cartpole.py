"""QCartPoleSwingUpSim -- Quanser cart-pole (pyrado/environments/pysim/quanser_cartpole.py:54-588), including the
fork's differentiable `step_diff_state` (:257-358) whose backward pass is the hand-written adjoint kernel."""
from typing import Optional

import numpy as np
import torch

import simurlacra_b200 as pyrado

from ... import _lib as L
from ... import engine as E
from ...spaces.box import BoxSpace
from ...tasks.desired_state import RadiallySymmDesStateTask
from ...tasks.reward_functions import ExpQuadrErrRewFcn
from .base import SimPyEnv

MAX_ACT_QCP = 6.0  # [V]


class QCartPoleSwingUpSim(SimPyEnv):
    name: str = "qcp-su"
    kind = L.QCP
    _param_names = ("gravity_const", "cart_mass", "rail_length", "motor_efficiency", "gear_efficiency", "gear_ratio",
                    "motor_inertia", "pinion_radius", "motor_resistance", "motor_back_emf", "pole_damping",
                    "combined_damping", "pole_mass", "pole_length", "cart_friction_coeff", "voltage_thold_neg",
                    "voltage_thold_pos")

    def __init__(self, dt: float, max_steps: int = pyrado.inf, task_args: Optional[dict] = None, long: bool = False,
                 simple_dynamics: bool = False, wild_init: str = "True", mass=None, **kwargs):
        """
        :param long: long (0.641 m) or short pole
        :param simple_dynamics: Quanser's simplified dynamics (no Coulomb friction, no dead zone)
        :param wild_init: the fork compares STRINGS: 'True' -> wide init box, 'False' -> narrow one, anything else
                          (including the bool True that sac_wgrad passes) -> the third box (quanser_cartpole.py:552-560)
        """
        self._long, self._mass = long, mass
        self._wild_init_arg = wild_init
        self._x_buffer = 0.15
        super().__init__(dt, max_steps, task_args, **kwargs)
        self._flags = L.F_SIMPLE_DYN if simple_dynamics else 0
        self._wild_init = 0 if wild_init == "True" else (1 if wild_init == "False" else 2)
        self.domain_param = self.get_nominal_domain_param(long=long, mass=mass)

    @classmethod
    def get_nominal_domain_param(cls, long: bool = False, mass=None) -> dict:  # :111-143
        m_pole, l_pole = (0.23, 0.641 / 2) if long else (0.127, 0.3365 / 2)
        if mass is not None:
            m_pole = mass
        return dict(gravity_const=9.81, cart_mass=0.58, rail_length=0.814, motor_efficiency=0.9, gear_efficiency=0.9,
                    gear_ratio=3.71, motor_inertia=3.9e-7, pinion_radius=6.35e-3, motor_resistance=2.6,
                    motor_back_emf=7.67e-3, pole_damping=0.0024, combined_damping=5.4, pole_mass=m_pole,
                    pole_length=l_pole, cart_friction_coeff=0.02, voltage_thold_neg=0.0, voltage_thold_pos=0.0)

    def _create_spaces(self):  # :88-95, :541-560
        l_rail = self._domain_param["rail_length"]
        max_obs = np.array([l_rail / 2.0, 1.0, 1.0, np.inf, np.inf])
        self._obs_space = BoxSpace(-max_obs, max_obs, labels=["x", "sin_theta", "cos_theta", "x_dot", "theta_dot"])
        self._act_space = BoxSpace(-MAX_ACT_QCP, MAX_ACT_QCP, shape=(1,), labels=["V"])
        max_state = np.array([l_rail / 2.0 - self._x_buffer, 4 * np.pi, 1 * l_rail, 20 * np.pi])
        w = getattr(self, "_wild_init_arg", "True")
        if w == "True":
            mi = np.array([0.25, np.pi, 0.8, np.pi])
        elif w == "False":
            mi = np.array([0.02, 2 / 180.0 * np.pi, 0.0, 1 / 180.0 * np.pi])
        else:
            mi = np.array([0.02, np.pi, 0.0, 1 / 180.0 * np.pi])
        self._state_space = BoxSpace(-max_state, max_state, labels=["x", "theta", "x_dot", "theta_dot"])
        if not self._custom_init_space:
            self._init_space = BoxSpace(-mi, mi, labels=["x", "theta", "x_dot", "theta_dot"])

    def _create_task(self, task_args: dict):  # :573-588
        state_des = np.asarray(task_args.get("state_des", np.array([0.0, np.pi, 0.0, 0.0])), dtype=np.float64)
        Q = task_args.get("Q", np.diag([3e-1, 5e-1, 5e-3, 1e-3]))
        R = task_args.get("R", np.diag([1e-3]))
        return RadiallySymmDesStateTask(self.spec, state_des, ExpQuadrErrRewFcn(Q, R), idcs=[1])

    # ---- differentiable step (quanser_cartpole.py:257-280): torch in, torch out, autograd through the adjoint kernel
    def step_diff_state(self, state: torch.Tensor, act: torch.Tensor, act_diff: torch.Tensor = None, th_ddot=None):
        """state [B, 4], act [B, 1] in volts (wrappers de-normalise before), th_ddot [B] carry or None.
        Returns next_state [B, 4], rew [B], done [B] (bool), {'th_ddot': [B]}; differentiable w.r.t. state, act,
        act_diff and th_ddot."""
        from ...sampling.diff import QcpDiffStep

        if act_diff is None:
            act_diff = torch.zeros_like(act)
        a_un = (act + act_diff).reshape(-1)
        if th_ddot is None:
            th_ddot = torch.zeros(state.shape[0], dtype=state.dtype, device=state.device)
        cfg, params = self._diff_cfg(state.dtype), self._diff_params(state.shape[0], state.dtype)
        ns, rew, done, thdd = QcpDiffStep.apply(cfg, params, state, a_un, th_ddot.reshape(-1))
        return ns, rew, done, {"th_ddot": thdd}

    def _diff_cfg(self, dtype):
        """C-ABI config of the differentiable step in the working dtype of the tensors."""
        return E.make_cfg(self.kind, self._dt, self._max_steps, dtype=dtype,
                          simple_dynamics=bool(self._flags & L.F_SIMPLE_DYN), wild_init=self._wild_init,
                          task=self._task_struct(), base_params=self.base_param_vector())

    def _diff_params(self, n: int, dtype):
        """Per-env parameter rows if the env holds one row per lane of the batch, else None (the base parameters)."""
        params = self._params if self._params is not None and self._params.shape[0] == n else None
        if params is not None and params.dtype != dtype:
            params = params.to(dtype)
        return params

    def step_diff(self, obs: torch.Tensor, act: torch.Tensor, th_ddot=None):
        """The observation-based differentiable step (quanser_cartpole.py:232-254 with `_step_dynamics_diff(state_based=
        False)`, :332-340): obs [B, 5] = [x, sin th, cos th, x_dot, th_dot]; theta is rebuilt from (sin, cos) with the
        fork's acos / asin average, the step itself is the RK4 kernel with its adjoint, and the NEXT OBSERVATION comes
        back.  act [B, 1] in volts (clamped inside; the reward sees the unclamped action)."""
        x, sin_th, cos_th, x_dot, th_dot = obs[:, 0], obs[:, 1], obs[:, 2], obs[:, 3], obs[:, 4]
        mask_cos, mask_sin = (cos_th > 0).to(obs.dtype), (sin_th > 0).to(obs.dtype)
        # the reference builds the masks with `.float()`, so its `np.pi * mask` term is a float32 product whatever the
        # dtype of the observation (:336-340): pi enters rounded to float32, reproduced here
        pi32 = float(np.float32(np.pi))
        th = ((mask_sin * 2 - 1) * torch.acos(cos_th) + (mask_cos * 2 - 1) * torch.asin(sin_th)
              + pi32 * (2 * mask_sin - 1) * (1 - mask_cos)) / 2
        state = torch.stack([x, th, x_dot, th_dot], dim=-1)
        ns, rew, done, info = self.step_diff_state(state, act, None, th_ddot)
        nobs = torch.stack([ns[:, 0], torch.sin(ns[:, 1]), torch.cos(ns[:, 1]), ns[:, 2], ns[:, 3]], dim=-1)
        return nobs, rew, done, info

    def _task_struct(self):
        t = L.Task()
        self._task.fill(t)
        return t


class QCartPoleStabSim(QCartPoleSwingUpSim):
    """Stabilise the pole in the upright position (pyrado/environments/pysim/quanser_cartpole.py:441-500): the cart-pole
    dynamics of the swing-up env with theta bounded to pi +- 15 deg (leaving it fails the episode), a narrow init
    space around the upright pose and a quadratic cost with a state- and time-dependent final cost."""

    name: str = "qcp-st"
    _pass_init_box = True

    def __init__(self, dt: float, max_steps: int = pyrado.inf, task_args: Optional[dict] = None, long: bool = True,
                 simple_dynamics: bool = True, **kwargs):
        self.stab_thold = 15 / 180.0 * np.pi  # [rad] failure threshold of the stabilisation task
        self.max_init_th_offset = 8 / 180.0 * np.pi  # [rad]
        super().__init__(dt, max_steps, task_args, long, simple_dynamics, wild_init="False", **kwargs)
        self._flags |= L.F_QCP_STAB

    def _create_spaces(self):  # :473-489
        super()._create_spaces()
        l_rail = self._domain_param["rail_length"]
        min_state = np.array([-l_rail / 2.0 + self._x_buffer, np.pi - self.stab_thold, -l_rail, -2 * np.pi])
        max_state = np.array([+l_rail / 2.0 - self._x_buffer, np.pi + self.stab_thold, +l_rail, +2 * np.pi])
        self._state_space = BoxSpace(min_state, max_state, labels=["x", "theta", "x_dot", "theta_dot"])
        if not self._custom_init_space:
            max_init = np.array([+0.02, np.pi + self.max_init_th_offset, +0.02, +5 / 180 * np.pi])
            min_init = np.array([-0.02, np.pi - self.max_init_th_offset, -0.02, -5 / 180 * np.pi])
            self._init_space = BoxSpace(min_init, max_init, labels=["x", "theta", "x_dot", "theta_dot"])

    def _create_task(self, task_args: dict):  # :491-500
        from ...tasks.final_reward import FinalRewMode, FinalRewTask
        from ...tasks.reward_functions import QuadrErrRewFcn

        state_des = np.asarray(task_args.get("state_des", np.array([0.0, np.pi, 0.0, 0.0])), dtype=np.float64)
        Q = task_args.get("Q", np.diag([5e-0, 1e1, 1e-2, 1e-2]))
        R = task_args.get("R", np.diag([1e-3]))
        return FinalRewTask(RadiallySymmDesStateTask(self.spec, state_des, QuadrErrRewFcn(Q, R), idcs=[1]),
                            mode=FinalRewMode(state_dependent=True, time_dependent=True))
