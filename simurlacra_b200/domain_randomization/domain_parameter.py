"""
DomainParam family (pyrado/domain_randomization/domain_parameter.py:46-312): Uniform, Normal and Bernoulli domain
parameters.  The reference draws from ``torch.distributions`` on the host generator; here every draw comes from the
device-side counter-based Philox stream (csrc/sre_core.cuh ``sample_entry``), keyed on
``(seed, sample index, draw counter, column)``, so that the same seed gives the same parameters for any batch split or
GPU count.  Samples are float32 like the reference's (torch default dtype), clamped and optionally rounded on device.
"""
import math
from typing import List, Optional, Union

import torch

import simurlacra_b200 as pyrado


class DomainParam:
    """Stores one domain parameter (a.k.a. physics parameter) and its distribution."""

    _distr_kind = None  # 'normal' | 'uniform' | 'bernoulli'

    def __init__(self, name: Union[str, List[str]], clip_lo: Optional[Union[int, float]] = -pyrado.inf,
                 clip_up: Optional[Union[int, float]] = pyrado.inf, roundint: bool = False):
        if not isinstance(name, (str, list)):
            raise pyrado.TypeErr(given=name, expected_type=(str, list))
        self.name = name
        self.mean = None
        self.clip_lo = clip_lo
        self.clip_up = clip_up
        self.roundint = roundint
        self._draws = 0  # Philox draw counter of stand-alone sample() calls

    def __eq__(self, other):
        if not isinstance(other, DomainParam):
            raise pyrado.TypeErr(given=other, expected_type=DomainParam)
        return all(getattr(self, fn) == getattr(other, fn, None) for fn in self.get_field_names())

    __hash__ = None

    def get_field_names(self) -> List[str]:
        return ["name", "clip_lo", "clip_up", "roundint"]

    def adapt(self, domain_distr_param: str, domain_distr_param_value):
        """Update one distribution parameter, e.g. mean or std (domain_parameter.py:87-102)."""
        if domain_distr_param not in self.get_field_names():
            raise pyrado.KeyErr(msg=f"The domain parameter {self.name} does not have a domain distribution parameter "
                                    f"called {domain_distr_param}!")
        setattr(self, domain_distr_param, domain_distr_param_value)
        self._validate()

    def _validate(self):
        pass

    def entry(self, param_index: int) -> dict:
        """This parameter as one row of the C-ABI distribution table (``sre_distr``)."""
        raise RuntimeError("Trying to sample a domain parameter without a specified distribution!")

    def sample(self, num_samples: int = 1, device=None) -> List[torch.Tensor]:
        """``num_samples`` draws as a list of 0-d float32 tensors (int32 when ``roundint``), domain_parameter.py:104-132."""
        if not isinstance(num_samples, int):
            raise pyrado.TypeErr(given=num_samples, expected_type=int)
        if num_samples <= 0:
            raise pyrado.ValueErr(given=num_samples, g_constraint="0")
        from .domain_randomizer import sample_table

        self._draws += 1
        col = sample_table([self.entry(0)], num_samples, episode=self._draws, device=device)[:, 0]
        if self.roundint:
            col = col.to(torch.int32)
        return list(col.cpu())


class UniformDomainParam(DomainParam):
    """Domain parameter sampled from U(mean - halfspan, mean + halfspan)"""

    _distr_kind = "uniform"

    def __init__(self, mean, halfspan, **kwargs):
        super().__init__(**kwargs)
        self.mean = mean
        self.halfspan = halfspan
        self._validate()

    def _validate(self):
        if not float(self.halfspan) > 0:  # torch.distributions.Uniform(validate_args=True) needs low < high
            raise ValueError(f"Uniform domain parameter {self.name}: halfspan must be positive, got {self.halfspan}")

    def get_field_names(self) -> List[str]:
        return ["name", "mean", "halfspan", "clip_lo", "clip_up", "roundint"]

    def entry(self, param_index: int) -> dict:
        return dict(param_index=param_index, kind="uniform", a=float(self.mean) - float(self.halfspan),
                    b=float(self.mean) + float(self.halfspan), clip_lo=self.clip_lo, clip_up=self.clip_up,
                    roundint=self.roundint)


class NormalDomainParam(DomainParam):
    """Domain parameter sampled from N(mean, std)"""

    _distr_kind = "normal"

    def __init__(self, mean, std, **kwargs):
        super().__init__(**kwargs)
        self.mean = mean
        self.std = std
        self._validate()

    def _validate(self):
        if not float(self.std) > 0 or not math.isfinite(float(self.mean)):
            raise ValueError(f"Normal domain parameter {self.name}: need finite mean and std > 0, got "
                             f"loc = {self.mean}, scale = {self.std}")

    def get_field_names(self) -> List[str]:
        return ["name", "mean", "std", "clip_lo", "clip_up", "roundint"]

    def entry(self, param_index: int) -> dict:
        return dict(param_index=param_index, kind="normal", a=float(self.mean), b=float(self.std),
                    clip_lo=self.clip_lo, clip_up=self.clip_up, roundint=self.roundint)


class BernoulliDomainParam(DomainParam):
    """Domain parameter taking val_1 with probability prob_1, else val_0"""

    _distr_kind = "bernoulli"

    def __init__(self, val_0, val_1, prob_1: float, **kwargs):
        kwargs.pop("mean", None)  # the reference's own example passes mean=None (default_randomizers.py:137)
        super().__init__(**kwargs)
        self.val_0 = val_0
        self.val_1 = val_1
        self.prob_1 = prob_1
        self._validate()

    def _validate(self):
        if not 0.0 <= float(self.prob_1) <= 1.0:
            raise ValueError(f"Bernoulli domain parameter {self.name}: probs = {self.prob_1} is not a probability")
        self.mean = self.val_0 * (1 - self.prob_1) + self.val_1 * self.prob_1

    def get_field_names(self) -> List[str]:
        return ["name", "mean", "val_0", "val_1", "prob_1", "clip_lo", "clip_up", "roundint"]

    def entry(self, param_index: int) -> dict:
        return dict(param_index=param_index, kind="bernoulli", a=float(self.val_0), b=float(self.val_1),
                    c=float(self.prob_1), clip_lo=self.clip_lo, clip_up=self.clip_up, roundint=self.roundint)
