"""LinearPolicy (pyrado/policies/feed_back/linear.py:41-110): inner product of nonlinear features of the observation
with the policy parameters, `nn.Linear(num_feat, act_dim, bias=False)`."""
from typing import Optional

import torch
import torch.nn as nn

import simurlacra_b200 as pyrado

from ... import _lib as L
from ... import engine as E
from ...utils.data_types import EnvSpec
from ..base import Policy
from ..features import FeatureStack
from ..initialization import init_param


class LinearPolicy(Policy):
    name: str = "lin"

    def __init__(self, spec: EnvSpec, feats: FeatureStack, init_param_kwargs: Optional[dict] = None,
                 use_cuda: bool = False):
        if not isinstance(feats, FeatureStack):
            raise pyrado.TypeErr(given=feats, expected_type=FeatureStack)
        super().__init__(spec, use_cuda)
        self._feats = feats
        self.num_active_feat = feats.get_num_feat(spec.obs_space.flat_dim)
        self.net = nn.Linear(self.num_active_feat, spec.act_space.flat_dim, bias=False)
        self.init_param(None, **(init_param_kwargs or dict()))
        self.to(self.device)

    @property
    def features(self) -> FeatureStack:
        return self._feats

    def init_param(self, init_values: torch.Tensor = None, **kwargs):
        if init_values is None:
            init_param(self.net, **kwargs)
        else:
            self.param_values = init_values

    def eval_feats(self, obs: torch.Tensor) -> torch.Tensor:
        return self._feats(obs)

    def forward(self, obs: torch.Tensor) -> torch.Tensor:
        obs = obs.to(device=self.device)
        batched = obs.ndimension() == 2
        act = self.net(self.eval_feats(obs))
        return act if batched else act.flatten()

    def engine_desc(self, device, param_sets=None, group_size: int = 0):
        codes = self._feats.engine_codes(self.env_spec.obs_space.flat_dim)
        if codes is None or len(codes) > L.MAX_FEATS:
            return None
        if sum(len(c[1]) for c in codes if isinstance(c, tuple)) > L.MAX_FEAT_PARAMS:
            return None  # e.g. hundreds of random Fourier features: evaluated as a torch module
        weights = self._engine_weights(device) if param_sets is None else param_sets
        return E.make_policy(L.POL_LINEAR, weights, feats=codes, group_size=group_size)
