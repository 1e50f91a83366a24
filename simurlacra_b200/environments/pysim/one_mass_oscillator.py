"""OneMassOscillatorSim (pyrado/environments/pysim/one_mass_oscillator.py:43-116)."""
import numpy as np

from ... import _lib as L
from ...spaces.box import BoxSpace
from ...tasks.desired_state import DesStateTask
from ...tasks.final_reward import FinalRewMode, FinalRewTask
from ...tasks.reward_functions import QuadrErrRewFcn
from .base import SimPyEnv


class OneMassOscillatorSim(SimPyEnv):
    name: str = "omo"
    kind = L.OMO
    _param_names = ("mass", "stiffness", "damping")

    @classmethod
    def get_nominal_domain_param(cls) -> dict:  # :81-87
        return dict(mass=1.0, stiffness=30.0, damping=0.5)

    def _create_spaces(self):  # :52-66 -- the action bound is the stiffness, i.e. it depends on the domain parameters
        k = self._domain_param["stiffness"]
        max_state = np.array([1.0, 10.0])
        self._state_space = BoxSpace(-max_state, max_state, labels=["x", "x_dot"])
        self._obs_space = self._state_space
        if not self._custom_init_space:
            self._init_space = BoxSpace(np.array([-0.75 * max_state[0], -0.01 * max_state[1]]),
                                        np.array([-0.65 * max_state[0], +0.01 * max_state[1]]), labels=["x", "x_dot"])
        self._act_space = BoxSpace(-np.array([max_state[0] * k]), np.array([max_state[0] * k]), labels=["F"])

    def _create_task(self, task_args: dict):  # :68-79
        state_des = np.asarray(task_args.get("state_des", np.zeros(2)), dtype=np.float64)
        Q = task_args.get("Q", np.diag([1e1, 1e-2]))
        R = task_args.get("R", np.diag([1e-6]))
        return FinalRewTask(DesStateTask(self.spec, state_des, QuadrErrRewFcn(Q, R)), factor=1e3,
                            mode=FinalRewMode(always_negative=True))
