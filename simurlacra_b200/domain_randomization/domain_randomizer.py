"""
DomainRandomizer (pyrado/domain_randomization/domain_randomizer.py:49-330).  ``randomize(n)`` draws ``n`` parameter
sets with ONE launch of the device-side Philox sampler (``sre_sample_params``) and keeps them as a ``[n, k]`` float32
device tensor; ``get_params`` hands them out in the reference's formats (list of dicts / dict of lists, 0-d float32
numpy arrays or torch tensors).  ``table()`` compiles the randomizer into the C-ABI distribution table that the
persistent rollout kernel samples from in-flight (one draw per env-episode, no host round trip).
"""
from copy import deepcopy
from typing import List, Mapping, Optional, Sequence, Union

import numpy as np
import torch
from tabulate import tabulate

import simurlacra_b200 as pyrado

from .. import _lib as L
from .. import engine as E
from .domain_parameter import BernoulliDomainParam, DomainParam, NormalDomainParam, UniformDomainParam

_SCRATCH_KIND = L.QBB  # the widest parameter vector (dp = 20) serves as scratch table for free-standing sampling


def _device(device=None) -> torch.device:
    if device is not None:
        return torch.device(device)
    if not torch.cuda.is_available():
        raise L.EngineError("domain parameter sampling runs on the GPU (Philox kernel); no CUDA device is available "
                            "and simurlacra_b200 has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def sample_table(entries: Sequence[dict], num_samples: int, episode: int = 0, seed: Optional[int] = None,
                 env_offset: int = 0, device=None) -> torch.Tensor:
    """Draw ``[num_samples, len(entries)]`` float32 values: sample j of entry i comes from the Philox block keyed
    (seed, env_offset + j, episode, i) -- the same key the rollout kernel uses for env j, episode ``episode``."""
    k = len(entries)
    dp = E.dims(_SCRATCH_KIND)["dp"]
    if k > dp:
        raise pyrado.ValueErr(given=k, le_constraint=f"{dp} randomized parameters per randomizer")
    dev = _device(device)
    ent = [dict(e, param_index=i) for i, e in enumerate(entries)]
    cfg = E.make_cfg(_SCRATCH_KIND, 0.01, 1, dtype=torch.float32)
    out = torch.zeros(num_samples, dp, dtype=torch.float32, device=dev)
    E.sample_params(cfg, E.make_distr_table(ent), out, pyrado.device_seed() if seed is None else seed, env_offset,
                    episode)
    return out[:, :k]


class DomainRandomizer:
    """Class for executing the domain randomization"""

    def __init__(self, *domain_params: DomainParam):
        self.domain_params: List[DomainParam] = []
        self.add_domain_params(*domain_params)
        self._samples = None       # [n, k] float32 device tensor of the last randomize() call
        self._sample_names = None
        self._draws = 0            # Philox "episode" of free-standing randomize() calls

    def __str__(self):
        order = ["name", "mean", "std", "halfspan", "val_0", "val_1", "prob_1", "clip_lo", "clip_up", "roundint"]
        present = {fn for dp in self.domain_params for fn in dp.get_field_names()}
        headers = [h for h in order if h in present]
        return tabulate([[getattr(dp, h, None) for h in headers] for dp in self.domain_params], headers=headers,
                        tablefmt="simple")

    def add_domain_params(self, *domain_params: DomainParam, dp_mapping: Optional[Mapping[int, str]] = None):
        for dp in domain_params:
            if not isinstance(dp, DomainParam):
                raise pyrado.TypeErr(given=dp, expected_type=DomainParam)
            self.domain_params.append(dp)
        if dp_mapping is not None:
            for _, value in sorted(dp_mapping.items()):
                if not isinstance(value, str):
                    raise pyrado.TypeErr(given=value, expected_type=str)
                self.domain_params.append(DomainParam(name=value))

    # ------------------------------------------------------------------ engine side
    def names(self) -> List[str]:
        return [dp.name for dp in self.domain_params]

    def table(self, param_names: Sequence[str]):
        """Compile into the C-ABI distribution table for an env whose parameter vector is ordered ``param_names``.
        Entries the env does not know (wrapper parameters) are skipped, like a dict update would ignore them."""
        ent = []
        for dp in self.domain_params:
            if dp.name in param_names:
                ent.append(dp.entry(list(param_names).index(dp.name)))
        return E.make_distr_table(ent)

    def sample_tensor(self, num_samples: int, *, episode: int = 0, seed: Optional[int] = None, env_offset: int = 0,
                      device=None) -> torch.Tensor:
        """``[num_samples, k]`` float32 device tensor, one column per domain parameter (engine extension)."""
        return sample_table([dp.entry(0) for dp in self.domain_params], num_samples, episode, seed, env_offset, device)

    # ------------------------------------------------------------------ reference API
    def randomize(self, num_samples: int, **key):
        """Draw ``num_samples`` parameter sets (domain_randomizer.py:123-157)."""
        if not isinstance(num_samples, int):
            raise pyrado.TypeErr(given=num_samples, expected_type=int)
        if num_samples <= 0:
            raise pyrado.ValueErr(given=num_samples, g_constraint="0")
        if "episode" not in key:
            self._draws += 1
            key["episode"] = self._draws
        self._samples = self.sample_tensor(num_samples, **key)
        self._sample_names = self.names()

    def _host_sets(self, dtype: str, n: int) -> List[dict]:
        host = self._samples[:n].cpu()
        rnd = [dp.roundint for dp in self.domain_params]
        sets = []
        for j in range(host.shape[0]):
            d = dict()
            for i, name in enumerate(self._sample_names):
                v = host[j, i].to(torch.int32) if rnd[i] else host[j, i].clone()
                d[name] = v.numpy() if dtype == "numpy" else v
            sets.append(d)
        return sets

    def get_params(self, num_samples: int = -1, fmt: str = "list", dtype: str = "numpy") -> Union[list, dict]:
        """The perturbed parameters in the reference's formats (domain_randomizer.py:159-227)."""
        if not isinstance(num_samples, int):
            raise pyrado.TypeErr(given=num_samples, expected_type=int)
        if num_samples <= -2 or num_samples == 0:
            raise pyrado.ValueErr(msg="The number of samples needs to be -1 or a positive integer!")
        if fmt.lower() not in ("list", "dict"):
            raise pyrado.ValueErr(given=fmt, eq_constraint="list or dict")
        if dtype.lower() not in ("numpy", "torch"):
            raise pyrado.ValueErr(given=dtype, eq_constraint="numpy or torch")
        if self._samples is None:
            raise RuntimeError(f"Something when wrong during get_params({num_samples}, {fmt}, {dtype})!")
        stored = self._samples.shape[0]
        if num_samples == 1 or stored == 1:
            one = self._host_sets(dtype, 1)[0]  # a single set comes back as one dict (or a list holding it)
            return [one] if fmt == "list" else one
        n = stored if num_samples == -1 else min(num_samples, stored)
        sets = self._host_sets(dtype, n)
        if fmt == "list":
            return sets
        return {name: [s[name] for s in sets] for name in self._sample_names}

    def adapt_one_distr_param(self, domain_param_name: str, domain_distr_param: str, domain_distr_param_value):
        for dp in self.domain_params:
            if dp.name == domain_param_name:
                if domain_distr_param in dp.get_field_names():
                    dp.adapt(domain_distr_param, domain_distr_param_value)
                else:
                    raise pyrado.KeyErr(msg=f"The domain parameter {dp.name} does not have a domain distribution "
                                            f"parameter called {domain_distr_param}!")

    def rescale_distr_param(self, param: str, scale: float):
        """Rescale e.g. std or halfspan of all distributions (domain_randomizer.py:309-330)."""
        if not scale >= 0:
            raise pyrado.ValueErr(given=scale, ge_constraint="0")
        for dp in self.domain_params:
            if hasattr(dp, param):
                setattr(dp, param, scale * getattr(dp, param))

    def copy(self) -> "DomainRandomizer":
        c = DomainRandomizer(*deepcopy(self.domain_params))
        return c
