"""BallOnBeamSim and BallOnBeamDiscSim (pyrado/environments/pysim/ball_on_beam.py:41-136, :139-162)."""
import numpy as np

from ... import _lib as L
from ...spaces.box import BoxSpace
from ...spaces.compound import CompoundSpace
from ...spaces.discrete import DiscreteSpace
from ...tasks.desired_state import DesStateTask
from ...tasks.reward_functions import ScaledExpQuadrErrRewFcn
from .base import SimPyEnv


class BallOnBeamSim(SimPyEnv):
    name: str = "bob"
    kind = L.BOB
    _param_names = ("gravity_const", "ball_mass", "ball_radius", "beam_mass", "beam_length", "beam_thickness",
                    "friction_coeff", "ang_offset")

    @classmethod
    def get_nominal_domain_param(cls) -> dict:  # :75-87
        return dict(gravity_const=9.81, ball_mass=0.5, ball_radius=0.1, beam_mass=3.0, beam_length=2.0,
                    beam_thickness=0.1, friction_coeff=0.05, ang_offset=0.0)

    def _create_spaces(self):  # :48-73
        l_beam, g = self._domain_param["beam_length"], self._domain_param["gravity_const"]
        max_state = np.array([l_beam / 2.0, np.pi / 4.0, 10.0, np.pi])
        max_act = np.array([l_beam / 2.0 * g * 3.0])
        lab = ["x", "alpha", "x_dot", "alpha_dot"]
        self._state_space = BoxSpace(-max_state, max_state, labels=lab)
        self._obs_space = self._state_space
        if not self._custom_init_space:
            lo = np.array([-0.8 * l_beam / 2.0, -5 / 180.0 * np.pi, -0.02 * max_state[2], -0.02 * max_state[3]])
            up = np.array([-0.7 * l_beam / 2.0, +5 / 180.0 * np.pi, +0.02 * max_state[2], +0.02 * max_state[3]])
            lo2, up2 = lo.copy(), up.copy()
            lo2[0], up2[0] = 0.7 * l_beam / 2.0, 0.8 * l_beam / 2.0
            self._init_space = CompoundSpace([BoxSpace(lo, up, labels=lab), BoxSpace(lo2, up2, labels=lab)])
        self._act_space = BoxSpace(-max_act, max_act, labels=["tau"])

    def _create_task(self, task_args: dict):  # :100-108
        state_des = np.asarray(task_args.get("state_des", np.zeros(4)), dtype=np.float64)
        Q = task_args.get("Q", np.diag([1e5, 1e3, 1e3, 1e2]))
        R = task_args.get("R", np.eye(self.spec.act_space.flat_dim))
        return DesStateTask(self.spec, state_des,
                            ScaledExpQuadrErrRewFcn(Q, R, self.state_space, self.act_space, min_rew=1e-4))


class BallOnBeamDiscSim(BallOnBeamSim):
    """Ball-on-beam simulation environment with discrete actions: every action is projected onto
    {-max_act, 0, +max_act} (the kernels' `SRE_F_DISC_ACT`)."""

    name: str = "bob-d"

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._flags |= L.F_DISC_ACT

    def _create_spaces(self):  # :153-162
        super()._create_spaces()
        min_act, max_act = self._act_space.bounds
        self._act_space = DiscreteSpace(np.linspace(min_act, max_act, num=3, endpoint=True), labels=["tau"])
