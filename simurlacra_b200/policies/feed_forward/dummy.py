"""IdlePolicy and DummyPolicy (pyrado/policies/feed_forward/dummy.py:38-90)."""
import torch

from ... import _lib as L
from ... import engine as E
from ...utils.data_types import EnvSpec
from ..base import Policy


class IdlePolicy(Policy):
    """The most simple policy which simply does nothing"""

    name: str = "idle"

    def __init__(self, spec: EnvSpec, use_cuda: bool = False):
        super().__init__(spec, use_cuda)

    def init_param(self, init_values: torch.Tensor = None, **kwargs):
        pass

    def forward(self, obs: torch.Tensor = None) -> torch.Tensor:
        shape = self._env_spec.act_space.shape
        if obs is not None and obs.ndim == 2:
            shape = (obs.shape[0], *shape)
        return torch.zeros(shape, dtype=torch.get_default_dtype(), device=obs.device if obs is not None else self.device)

    def engine_desc(self, device):
        return E.make_policy(L.POL_IDLE)


class DummyPolicy(Policy):
    """Simple policy which samples random values form the action space (host generator, like the reference)"""

    name: str = "dummy"

    def __init__(self, spec: EnvSpec, use_cuda: bool = False):
        super().__init__(spec, use_cuda)

    def init_param(self, init_values: torch.Tensor = None, **kwargs):
        pass

    def forward(self, obs: torch.Tensor = None) -> torch.Tensor:
        sp = self.env_spec.act_space
        if obs is None or obs.ndim == 1:
            act = torch.from_numpy(sp.sample_uniform())
        else:
            act = torch.stack([torch.from_numpy(sp.sample_uniform()) for _ in range(obs.shape[0])], dim=0)
        return act.to(dtype=torch.get_default_dtype(), device=obs.device if obs is not None else self.device)
