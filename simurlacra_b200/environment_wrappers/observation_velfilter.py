"""
ObsVelFiltWrapper (pyrado/environment_wrappers/observation_velfilter.py:41-145): the velocity entries of the
observation are replaced by a discretised derivative filter num(s) / den(s) (default 50 s / (s + 50), like the filter
on the Quanser platforms) applied to the position entries of the STATE.  The filter starts in steady state for the
initial position.  One env (ndarray) or a batch (device tensors; the filter state is [order, num_envs, n_pos]).
"""
from typing import List, Optional, Tuple

import numpy as np
import torch
from scipy import signal

import simurlacra_b200 as pyrado

from ..environments.sim_base import SimEnv
from .base import EnvWrapperObs, _call_reset
from .utils import inner_env


class ObsVelFiltWrapper(EnvWrapperObs):
    """Environment wrapper which computes the velocities from the states given a linear filter"""

    def __init__(self, wrapped_env, mask_pos: Optional[List] = None, idcs_pos: Optional[List] = None,
                 mask_vel: Optional[List] = None, idcs_vel: Optional[List] = None, num: Optional[Tuple] = (50, 0),
                 den: Optional[Tuple] = (1, 50)):
        if not isinstance(inner_env(wrapped_env), SimEnv):
            raise pyrado.TypeErr(given=inner_env(wrapped_env), expected_type=SimEnv)
        super().__init__(wrapped_env)
        self._ctor = dict(mask_pos=mask_pos, idcs_pos=idcs_pos, mask_vel=mask_vel, idcs_vel=idcs_vel, num=num, den=den)
        if mask_pos is not None:
            self.mask_pos = np.array(mask_pos, dtype=bool)
            if not self.mask_pos.shape == wrapped_env.state_space.shape:
                raise pyrado.ShapeErr(given=mask_pos, expected_match=wrapped_env.state_space)
        else:
            if idcs_pos is None:
                raise pyrado.ValueErr(msg="Either mask or indices must be specified!")
            self.mask_pos = wrapped_env.state_space.create_mask(idcs_pos)
        if mask_vel is not None:
            self.mask_vel = np.array(mask_vel, dtype=bool)
            if not self.mask_vel.shape == wrapped_env.obs_space.shape:
                raise pyrado.ShapeErr(given=mask_vel, expected_match=wrapped_env.obs_space)
        else:
            if idcs_vel is None:
                raise pyrado.ValueErr(msg="Either mask or indices must be specified!")
            self.mask_vel = wrapped_env.obs_space.create_mask(idcs_vel)
        if int(self.mask_pos.sum()) != int(self.mask_vel.sum()):
            raise pyrado.ShapeErr(msg="As many velocity entries as position entries are needed!")
        # coefficient design on the host (a handful of numbers); float32 like the reference (:102-104)
        filt = signal.cont2discrete((num, den), dt=wrapped_env.dt)
        self.b = filt[0].ravel().astype(np.float32)
        self.a = filt[1].astype(np.float32)
        n = max(len(self.a), len(self.b))
        self._b = np.concatenate([self.b, np.zeros(n - len(self.b))]).astype(np.float64) / float(self.a[0])
        self._a = np.concatenate([self.a, np.zeros(n - len(self.a))]).astype(np.float64) / float(self.a[0])
        self.z = np.zeros((n - 1, int(self.mask_pos.sum())), dtype=np.float32)

    def _fuse_rollout(self, fusion, below) -> bool:
        return True  # the persistent rollout kernel keeps this wrapper's state per lane (sampling/rollout.py::_compile_step_wrappers)

    def _ctor_kwargs(self) -> dict:
        return dict(self._ctor)

    def _pos(self, state):
        if isinstance(state, torch.Tensor):
            return state[..., torch.as_tensor(np.flatnonzero(self.mask_pos), device=state.device)]
        return state[..., self.mask_pos]

    def init_filter(self, init_state):
        """Start the filter in steady state for the position part of `init_state` (:107-117)."""
        if not isinstance(init_state, (np.ndarray, torch.Tensor)):
            raise pyrado.TypeErr(given=init_state, expected_type=np.ndarray)
        zi = signal.lfilter_zi(self.b, self.a)
        pos = self._pos(init_state)
        if isinstance(pos, torch.Tensor):
            self.z = torch.tensor(zi.tolist(), dtype=pos.dtype, device=pos.device).reshape(-1, *([1] * pos.dim())) * pos
        else:
            self.z = np.outer(zi, pos)

    def _filter(self, x):
        """One sample through the transposed direct form II (what scipy.signal.lfilter runs): y = b0 x + z0,
        z_i = z_{i+1} + b_{i+1} x - a_{i+1} y."""
        z = self.z
        if isinstance(x, torch.Tensor) and not isinstance(z, torch.Tensor):
            z = torch.as_tensor(np.asarray(z, dtype=np.float64), dtype=x.dtype, device=x.device)
            z = z.reshape(z.shape[0], *([1] * (x.dim() - 1)), z.shape[-1]).expand(z.shape[0], *x.shape).clone()
        y = self._b[0] * x + z[0]
        new = []
        for i in range(len(self._b) - 1):
            nxt = z[i + 1] if i + 1 < len(self._b) - 1 else 0.0
            new.append(nxt + self._b[i + 1] * x - self._a[i + 1] * y)
        self.z = torch.stack(new) if isinstance(x, torch.Tensor) else np.stack(new)
        return y

    def _process_obs(self, obs):
        vel = self._filter(self._pos(self.state))
        if isinstance(obs, torch.Tensor):
            obs = obs.clone()
            obs[..., torch.as_tensor(np.flatnonzero(self.mask_vel), device=obs.device)] = vel.to(obs.dtype)
        else:
            obs = np.array(obs, copy=True)
            obs[..., self.mask_vel] = vel
        return obs

    def _reset_chain(self, init_state, domain_param, fusion):
        if domain_param is not None:
            self._get_wrapper_domain_param(domain_param)
        obs = _call_reset(self._wrapped_env, init_state, domain_param, None)
        self.init_filter(self.state)
        return self._process_obs(obs)
