from .base import Space  # noqa: F401
from .box import BoxSpace, InfBoxSpace  # noqa: F401
from .compound import CompoundSpace  # noqa: F401
from .discrete import DiscreteSpace  # noqa: F401
from .polar import Polar2DPosVelSpace  # noqa: F401
from .singular import SingularStateSpace  # noqa: F401
