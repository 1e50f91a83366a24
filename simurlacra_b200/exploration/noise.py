"""
Additive exploration noise modules: ``DiagNormalNoise`` (pyrado/exploration/normal_noise.py:40-172) and ``UniformNoise``
(pyrado/exploration/uniform_noise.py:38-177).  They stay torch ``nn.Module``s (the learner optimises ``log_std`` /
``log_halfspan`` with autograd); for sampling, their current scale and mean are copied into the C-ABI policy descriptor
(``engine_noise``) and the rollout kernel draws the noise itself from the Philox action-noise stream.
"""
import math
from typing import Optional, Union
from warnings import warn

import torch
import torch.nn as nn
from torch.distributions import Normal, Uniform

import simurlacra_b200 as pyrado

from .. import _lib as L


class _AdditiveNoise(nn.Module):
    """Shared plumbing: a log-transformed positive scale vector, an optional learnable mean, device bookkeeping."""

    _scale_name = "scale"
    engine_kind = L.ACT_NOISE_NONE

    def __init__(self, noise_dim, scale_init, scale_min, train_mean: bool, learnable: bool, use_cuda: bool):
        n = self._scale_name
        if not isinstance(scale_init, (float, int, torch.Tensor)):
            raise pyrado.TypeErr(given=scale_init, expected_type=[float, int, torch.Tensor])
        if isinstance(scale_init, torch.Tensor):
            dim = tuple(noise_dim) if isinstance(noise_dim, (tuple, list, torch.Size)) else (int(noise_dim),)
            if tuple(scale_init.shape) != dim:
                raise pyrado.ShapeErr(given=scale_init, expected_match=torch.empty(dim))
            if not bool((scale_init > 0).all()):
                raise pyrado.ValueErr(given=scale_init, g_constraint="0")
        elif not scale_init > 0:
            raise pyrado.ValueErr(given=scale_init, g_constraint="0")
        if not isinstance(scale_min, (float, torch.Tensor)):
            raise pyrado.TypeErr(given=scale_min, expected_type=[float, torch.Tensor])
        if isinstance(scale_min, torch.Tensor):
            if not bool((scale_min > 0).all()):
                raise pyrado.ValueErr(given=scale_min, g_constraint="0")
        elif not scale_min > 0:
            raise pyrado.ValueErr(given=scale_min, g_constraint="0")
        super().__init__()
        if use_cuda and not torch.cuda.is_available():
            warn("Tried to run on CUDA, but it is not available. Falling back to CPU.")
        self._device = "cuda" if use_cuda and torch.cuda.is_available() else "cpu"
        if learnable:
            setattr(self, "log_" + n, nn.Parameter(torch.empty(noise_dim), requires_grad=True))
            self.mean = nn.Parameter(torch.empty(noise_dim), requires_grad=True) if train_mean else None
        else:
            setattr(self, "log_" + n, torch.empty(noise_dim))
            self.mean = None
        dt = torch.get_default_dtype()
        init = torch.tensor(float(scale_init), dtype=dt) if isinstance(scale_init, (float, int)) else scale_init.to(dt)
        setattr(self, f"log_{n}_init", torch.log(init).to(self._device))
        setattr(self, n + "_min", (torch.tensor(scale_min, dtype=dt) if isinstance(scale_min, float) else scale_min).to(self._device))
        self.reset_expl_params()
        self.to(self._device)

    @property
    def device(self) -> str:
        """The device (CPU or GPU) on which the module is stored"""
        return self._device

    @property
    def _log_scale(self) -> torch.Tensor:
        return getattr(self, "log_" + self._scale_name)

    def _set_scale(self, value: torch.Tensor):
        self._log_scale.data = torch.log(torch.max(value, getattr(self, self._scale_name + "_min")))

    def reset_expl_params(self):
        """Reset all parameters of the exploration strategy."""
        if self.mean is not None:
            self.mean.data.zero_()
        self._log_scale.data.copy_(getattr(self, f"log_{self._scale_name}_init"))

    def _adapt(self, mean: Optional[torch.Tensor], scale, what: str):
        if not (isinstance(mean, torch.Tensor) or mean is None):
            raise pyrado.TypeErr(given=mean, expected_type=torch.Tensor)
        if not (isinstance(scale, torch.Tensor) and bool((scale >= 0).all()) or scale is None):
            raise pyrado.TypeErr(msg=f"The {what} must be a Tensor with all elements > 0 or None, but it is {scale}!")
        if mean is not None:
            assert self.mean is not None, "Can not change fixed zero mean!"
            if not mean.shape == self.mean.shape:
                raise pyrado.ShapeErr(given=mean, expected_match=self.mean)
            self.mean.data = mean
        if scale is not None:
            if not scale.shape == self._log_scale.shape:
                raise pyrado.ShapeErr(given=scale, expected_match=self._log_scale)
            self._set_scale(scale)

    def _loc(self, value: torch.Tensor) -> torch.Tensor:
        return value if self.mean is None else value + self.mean

    # ------------------------------------------------------------------ engine side
    def engine_noise(self):
        """(kind, scale, mean) for ``engine.set_act_noise``: the current float32 values, read once per sample()."""
        with torch.no_grad():
            scale = self._engine_scale().detach().to("cpu", torch.float32).reshape(-1).tolist()
            mean = [0.0] * len(scale) if self.mean is None else self.mean.detach().to("cpu", torch.float32).reshape(-1).tolist()
        return self.engine_kind, scale, mean


class DiagNormalNoise(_AdditiveNoise):
    """Module for learnable additive Gaussian noise with a diagonal covariance matrix"""

    _scale_name = "std"
    engine_kind = L.ACT_NOISE_NORMAL

    def __init__(self, noise_dim, std_init, std_min: Union[float, torch.Tensor] = 1e-3, train_mean: bool = False,
                 learnable: bool = True, use_cuda: bool = False):
        super().__init__(noise_dim, std_init, std_min, train_mean, learnable, use_cuda)

    @property
    def std(self) -> torch.Tensor:
        """The untransformed standard deviation (normal_noise.py:113-115: no lower bound on read)"""
        return torch.exp(self.log_std)

    @std.setter
    def std(self, std: torch.Tensor):
        self._set_scale(std)

    def _engine_scale(self):
        return self.std

    def adapt(self, mean: Optional[torch.Tensor] = None, std: Union[torch.Tensor, float, None] = None):
        """Adapt mean and standard deviation; ``None`` leaves a parameter unchanged (normal_noise.py:132-153)."""
        self._adapt(mean, std, "std")

    def forward(self, value: torch.Tensor) -> Normal:
        """The noise distribution around a noise-free value"""
        return Normal(self._loc(value), self.std)

    def get_entropy(self) -> torch.Tensor:
        """Entropy of the diagonal normal (independent of the mean)"""
        return 0.5 + 0.5 * math.log(2 * math.pi) + 0.5 * torch.log(torch.prod(torch.pow(self.std, 2)))


class UniformNoise(_AdditiveNoise):
    """Module for learnable additive uniform noise"""

    _scale_name = "halfspan"
    engine_kind = L.ACT_NOISE_UNIFORM

    def __init__(self, use_cuda: bool, noise_dim, halfspan_init, halfspan_min: Union[float, torch.Tensor] = 0.01,
                 train_mean: bool = False, learnable: bool = True):
        super().__init__(noise_dim, halfspan_init, halfspan_min, train_mean, learnable, use_cuda)

    @property
    def halfspan(self) -> torch.Tensor:
        """The untransformed half interval, bounded from below (uniform_noise.py:115-118)"""
        return torch.max(torch.exp(self.log_halfspan), self.halfspan_min)

    @halfspan.setter
    def halfspan(self, halfspan: torch.Tensor):
        self._set_scale(halfspan)

    def _engine_scale(self):
        return self.halfspan

    def adapt(self, mean: Optional[torch.Tensor] = None, halfspan: Union[torch.Tensor, float, None] = None):
        """Adapt mean and half interval; ``None`` leaves a parameter unchanged (uniform_noise.py:134-158)."""
        self._adapt(mean, halfspan, "halfspan")

    def forward(self, value: torch.Tensor) -> Uniform:
        loc = self._loc(value)
        return Uniform(low=loc - self.halfspan, high=loc + self.halfspan)

    def get_entropy(self) -> torch.Tensor:
        """Entropy of the uniform distribution (independent of the mean)"""
        return torch.log(2 * self.halfspan)
