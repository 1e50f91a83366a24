"""
ActNormWrapper, ObsActCatWrapper and the straight-through clamp (pyrado/environment_wrappers/action_normalization.py).
The de-normalisation ``lb + (a + 1)(ub - lb)/2`` (:69-75) is fused into the step / rollout kernels (``cfg.act_norm``);
it only runs on the tensors themselves when a non-fusable action wrapper sits between this wrapper and the simulation.
"""
import numpy as np
import torch

from ..spaces.box import BoxSpace
from .base import EnvWrapper, EnvWrapperAct, inner_env_of


class DifferentiableClamp(torch.autograd.Function):
    """``torch.clamp`` in the forward pass, identity in the backward pass (:35-54)."""

    @staticmethod
    def forward(ctx, input, min, max):
        return input.clamp(min=min, max=max)

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None


def dclamp(input, min, max):
    """Like ``torch.clamp``, but with a constant 1-gradient (:57-63)."""
    return DifferentiableClamp.apply(input, min, max)


class ActNormWrapper(EnvWrapperAct):
    """Environment wrapper which normalizes the action space, such that all action values are in range [-1, 1]."""

    _fusable = True

    def _fuse(self, fusion, below_alters_act, below_alters_obs) -> bool:
        from .action_noise import GaussianActNoiseWrapper

        # a fused GaussianActNoiseWrapper may sit below (noise in env units); anything else runs on the data
        if any(w._alters_act and type(w) is not GaussianActNoiseWrapper for w in self._below()):
            return False
        fusion.act_norm = True
        return True

    def _fuse_rollout(self, fusion, below) -> bool:
        from .action_delay import ActDelayWrapper
        from .action_noise import GaussianActNoiseWrapper
        from .downsampling import DownsamplingWrapper

        # the persistent kernel also runs an action delay (in env units) and the downsampling hold below the normalisation
        if any(w._alters_act and type(w) not in (GaussianActNoiseWrapper, ActDelayWrapper, DownsamplingWrapper) for w in below):
            return False
        fusion.act_norm = True
        return True

    def _process_act(self, act):
        lb, ub = self.wrapped_env.act_space.bounds
        if isinstance(act, torch.Tensor):
            return self._process_act_tensor(act)
        return lb + (act + 1) * (ub - lb) / 2  # may leave the action space; the environment clips (base.py:223)

    def _process_act_tensor(self, act: torch.Tensor) -> torch.Tensor:
        lb, ub = self.wrapped_env.act_space.bounds
        lb, ub = torch.as_tensor(lb).to(act).unsqueeze(0), torch.as_tensor(ub).to(act).unsqueeze(0)
        return lb + (act + 1) * (ub - lb) / 2

    def _process_act_space(self, space: BoxSpace) -> BoxSpace:
        if not isinstance(space, BoxSpace):
            raise NotImplementedError("Only implemented ActNormWrapper._process_act_space() for BoxSpace!")
        return BoxSpace(-np.ones(space.shape), np.ones(space.shape), labels=space.labels)


class ObsActCatWrapper(EnvWrapper):
    """The fork's wrapper around ActNorm(QCartPole) that sac_wgrad uses (:91-200): observations are the inner env's
    STATES (obs_space is the state space), `step_diff_state` passes a zero action offset down the chain."""

    _alters_obs = True
    _fusable = False

    def _fuse(self, fusion, below_alters_act, below_alters_obs) -> bool:
        return False

    @property
    def obs_space(self) -> BoxSpace:
        return inner_env_of(self).state_space

    def _state_obs(self):
        s = inner_env_of(self).state
        return s.clone() if isinstance(s, torch.Tensor) else np.array(s, copy=True)

    def _reset_chain(self, init_state, domain_param, fusion):
        super()._reset_chain(init_state, domain_param, fusion)
        self.act_prev = np.array([0.0])
        return self._state_obs()  # the fork's cart-pole reset hands out the state (quanser_cartpole.py:101-106)

    def _step_chain(self, act, fusion) -> tuple:
        _, rew, done, info = super()._step_chain(act, fusion)
        a = act.detach().cpu().numpy() if isinstance(act, torch.Tensor) else np.asarray(act)
        self.act_prev = np.maximum(np.minimum(a, 1), -1)
        return self._state_obs(), rew, done, info

    def step_diff_state(self, obs, act, th_ddot=None):
        return self._wrapped_env.step_diff_state(obs, act, act * 0, th_ddot)

    def step_diff(self, obs, act, th_ddot=None):
        """:174-179: the last observation column is the previous (normalised) action; the env is stepped with
        `act + previous action`, and the new "previous action" is that sum through the straight-through clamp."""
        obs_state = obs[:, :-1]
        obsn, rew, done, info = self._wrapped_env.step_diff(obs_state, act + obs[:, -1:], th_ddot)
        act_prev = dclamp(act + obs[:, -1:], torch.tensor(-1.0).to(act), torch.tensor(1.0).to(act))
        info = dict(info, act_prev=act_prev)
        return obsn, rew, done, info
