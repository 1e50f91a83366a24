from .base import Env  # noqa: F401
from .sim_base import SimEnv  # noqa: F401
