"""EnvSpec and RenderMode (pyrado/utils/data_types.py:45-60)."""
from typing import NamedTuple, Optional


class EnvSpec(NamedTuple):
    """The specification of an environment's input and output space"""

    obs_space: object
    act_space: object
    state_space: Optional[object] = None


class RenderMode(NamedTuple):
    """Rendering is unavailable on this box (no graphics stack); the type is kept so that call sites type-check."""

    text: bool = False
    video: bool = False
    render: bool = False
