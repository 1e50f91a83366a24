"""Policies on the pysim sampling path (pyrado/policies): FNN, Linear-on-features, Idle, Dummy."""
from .base import Policy  # noqa: F401
from .features import FeatureStack  # noqa: F401
from .feed_back.fnn import FNN, FNNPolicy  # noqa: F401
from .feed_back.linear import LinearPolicy  # noqa: F401
from .feed_forward.dummy import DummyPolicy, IdlePolicy  # noqa: F401
