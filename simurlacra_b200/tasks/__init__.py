from .base import Task, TaskWrapper  # noqa: F401
from .desired_state import DesStateTask, RadiallySymmDesStateTask  # noqa: F401
from .final_reward import FinalRewMode, FinalRewTask  # noqa: F401
from .reward_functions import ExpQuadrErrRewFcn, QuadrErrRewFcn, RewFcn, ScaledExpQuadrErrRewFcn  # noqa: F401
from .utils import never_succeeded  # noqa: F401
