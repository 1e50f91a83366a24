"""SingularStateSpace (pyrado/spaces/singular.py:36-52): a box that contains exactly one state."""
import numpy as np

from .box import BoxSpace


class SingularStateSpace(BoxSpace):
    def __init__(self, fixed_state: np.ndarray, labels=None):
        fixed_state = np.asarray(fixed_state, dtype=np.float64)
        super().__init__(fixed_state, fixed_state, labels=labels)
        self._fixed_state = fixed_state.copy()

    def sample_uniform(self, concrete_inf: float = 1e6) -> np.ndarray:
        return self._fixed_state.copy()
