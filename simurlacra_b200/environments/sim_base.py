"""SimEnv interface (pyrado/environments/sim_base.py:40-123)."""
from abc import abstractmethod
from typing import List

from .base import Env


class SimEnv(Env):
    @property
    @abstractmethod
    def init_space(self):
        raise NotImplementedError

    @classmethod
    @abstractmethod
    def get_nominal_domain_param(cls) -> dict:
        raise NotImplementedError

    @property
    def supported_domain_param(self) -> List[str]:
        return list(self.get_nominal_domain_param().keys())

    @property
    @abstractmethod
    def domain_param(self) -> dict:
        raise NotImplementedError

    @domain_param.setter
    @abstractmethod
    def domain_param(self, param: dict):
        raise NotImplementedError
