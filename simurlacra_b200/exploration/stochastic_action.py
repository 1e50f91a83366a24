"""
Step-based exploration strategies (pyrado/exploration/stochastic_action.py:42-235): a policy wrapper that samples the
action from a distribution around the wrapped policy's output.

For the samplers the wrapper compiles to the wrapped policy's C-ABI descriptor plus the noise kind / scale / mean
(``engine_desc``), and the rollout kernel adds the noise itself (one Philox block per env-step, stream "action
noise"), so explorative sampling for PPO-style algorithms costs no host round trip.  ``forward`` (torch) is what the
step-by-step path and the learner use; ``evaluate`` re-evaluates a rollout and returns the action distribution with
gradient information, as in the reference.
"""
from abc import ABC, abstractmethod

import torch
from torch.distributions import Distribution

from .. import _lib as L
from .. import engine as E
from ..policies.base import Policy
from .noise import DiagNormalNoise, UniformNoise


class StochasticActionExplStrat(Policy, ABC):
    """Explore by sampling actions from a distribution."""

    def __init__(self, policy: Policy):
        super().__init__(policy.env_spec, use_cuda=policy.device != "cpu")
        self.policy = policy

    @property
    def is_recurrent(self) -> bool:
        return self.policy.is_recurrent

    def init_hidden(self, batch_size: int = None) -> torch.Tensor:
        return self.policy.init_hidden(batch_size)

    def init_param(self, init_values: torch.Tensor = None, **kwargs):
        self.policy.init_param(init_values, **kwargs)

    def reset(self, **kwargs):
        self.policy.reset(**kwargs)

    def forward(self, obs: torch.Tensor, *extra) -> torch.Tensor:
        # rsample: reparametrised, so that gradients can flow into the noise parameters (stochastic_action.py:81-82)
        out = self.policy(obs, *extra)
        if isinstance(out, tuple):  # recurrent / two-headed policies return (act, other...)
            return (self.action_dist_at(out[0]).rsample(),) + tuple(out[1:])
        return self.action_dist_at(out).rsample()

    def evaluate(self, rollout, hidden_states_name: str = "hidden_states") -> Distribution:
        """Re-evaluate a rollout with the wrapped policy: the action distribution with gradient data (:96-109)."""
        acts = self.policy.evaluate(rollout, hidden_states_name)
        if isinstance(acts, tuple):
            acts = acts[0]
        return self.action_dist_at(acts)

    @abstractmethod
    def action_dist_at(self, policy_output: torch.Tensor) -> Distribution:
        """The action distribution for a noise-free output of the wrapped policy"""
        raise NotImplementedError

    # ------------------------------------------------------------------ engine side
    @property
    def engine_precision(self):
        """Arithmetic of the fused hidden layer: the wrapped policy's setting"""
        pol = self.__dict__.get("_modules", {}).get("policy")
        return pol.engine_precision if pol is not None else L.PREC_TF32X3

    @engine_precision.setter
    def engine_precision(self, value):
        pol = self.__dict__.get("_modules", {}).get("policy")
        if pol is not None:
            pol.engine_precision = value

    def engine_desc(self, device, param_sets=None, group_size: int = 0):
        desc = self.policy.engine_desc(device, param_sets, group_size)
        if desc is None:
            return None
        kind, scale, mean = self.noise.engine_noise()
        return E.set_act_noise(desc, kind, scale, mean)

    def forward_engine(self, obs: torch.Tensor) -> torch.Tensor:
        """Noise-free batch forward of the wrapped policy (the stand-alone policy kernel draws no noise)."""
        return self.policy.forward_engine(obs)


class NormalActNoiseExplStrat(StochasticActionExplStrat):
    """Exploration strategy which adds Gaussian noise to the continuous policy actions"""

    def __init__(self, policy: Policy, std_init, std_min=1e-3, train_mean: bool = False, learnable: bool = True):
        super().__init__(policy)
        self._noise = DiagNormalNoise(noise_dim=policy.env_spec.act_space.flat_dim, std_init=std_init, std_min=std_min,
                                      train_mean=train_mean, learnable=learnable, use_cuda=policy.device != "cpu")

    def reset_expl_params(self, *args, **kwargs):
        self._noise.reset_expl_params(*args, **kwargs)

    def get_entropy(self, *args, **kwargs):
        return self._noise.get_entropy(*args, **kwargs)

    @property
    def std(self):
        return self._noise.std

    @std.setter
    def std(self, value):
        self._noise.std = value

    @property
    def mean(self):
        return self._noise.mean

    @mean.setter
    def mean(self, value):
        self._noise.mean = value

    @property
    def noise(self) -> DiagNormalNoise:
        """The exploration noise module"""
        return self._noise

    def action_dist_at(self, policy_output: torch.Tensor) -> Distribution:
        return self._noise(policy_output)


class UniformActNoiseExplStrat(StochasticActionExplStrat):
    """Exploration strategy which adds uniform noise to the continuous policy actions"""

    def __init__(self, policy: Policy, halfspan_init, halfspan_min=0.01, train_mean: bool = False, learnable: bool = True):
        super().__init__(policy)
        self._noise = UniformNoise(use_cuda=policy.device != "cpu", noise_dim=policy.env_spec.act_space.flat_dim,
                                   halfspan_init=halfspan_init, halfspan_min=halfspan_min, train_mean=train_mean,
                                   learnable=learnable)

    @property
    def noise(self) -> UniformNoise:
        return self._noise

    def action_dist_at(self, policy_output: torch.Tensor) -> Distribution:
        return self._noise(policy_output)

    def reset_expl_params(self, *args, **kwargs):
        self._noise.reset_expl_params(*args, **kwargs)

    def get_entropy(self, *args, **kwargs):
        return self._noise.get_entropy(*args, **kwargs)

    @property
    def halfspan(self):
        return self._noise.halfspan

    @halfspan.setter
    def halfspan(self, val):
        self._noise.halfspan = val
