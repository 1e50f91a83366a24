"""SamplerBase (pyrado/sampling/sampler.py:38-100): a sampler guarantees a minimum number of rollouts and / or steps."""
from abc import ABC, abstractmethod
from typing import List, Optional


class SamplerBase(ABC):
    def __init__(self, *, min_rollouts: int = None, min_steps: int = None):
        self.min_rollouts = None
        self.min_steps = None
        self.set_min_count(min_rollouts, min_steps)

    def set_min_count(self, min_rollouts: int = None, min_steps: int = None):
        assert min_rollouts is not None or min_steps is not None, "At least one limit must be given"
        self.min_rollouts = min_rollouts
        self.min_steps = min_steps

    @abstractmethod
    def reinit(self, env=None, policy=None):
        raise NotImplementedError

    @abstractmethod
    def sample(self) -> List:
        raise NotImplementedError
