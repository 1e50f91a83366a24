from .ball_on_beam import BallOnBeamSim  # noqa: F401
from .base import SimPyEnv  # noqa: F401
from .one_mass_oscillator import OneMassOscillatorSim  # noqa: F401
from .pendulum import PendulumSim  # noqa: F401
from .quanser_ball_balancer import QBallBalancerSim  # noqa: F401
from .quanser_cartpole import QCartPoleStabSim, QCartPoleSwingUpSim  # noqa: F401
from .quanser_qube import QQubeStabSim, QQubeSwingUpSim  # noqa: F401
