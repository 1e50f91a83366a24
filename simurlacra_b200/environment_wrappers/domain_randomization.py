"""
DomainRandWrapper, DomainRandWrapperLive, DomainRandWrapperBuffer, MetaDomainRandWrapper
(pyrado/environment_wrappers/domain_randomization.py:47-300).

The reference's flow is `randomizer.randomize(1)` -> `get_params(fmt="dict", dtype="numpy")` ->
`reset(domain_param=...)`.  Around a `SimPyEnv` one launch of the Philox sampler writes a fresh parameter set per env
straight into the env's `[num_envs, dp]` device tensor, keyed (seed, global env index, episode) -- the key the
persistent rollout kernel uses when it samples the same randomizer in-flight, so stepping through the wrapper and
the fused rollout see the same parameters.  Around any other SimEnv the reference's dict flow is used.
"""
from random import randint
from typing import List, Mapping, Optional, Tuple, Union

import numpy as np

import simurlacra_b200 as pyrado

from ..domain_randomization.domain_randomizer import DomainRandomizer
from ..environments.base import Env
from ..environments.pysim.base import SimPyEnv
from ..environments.sim_base import SimEnv
from .base import EnvWrapper
from .utils import all_envs, inner_env, remove_env, typed_env


class DomainRandWrapper(EnvWrapper):
    """Base class for environment wrappers which call a `DomainRandomizer` to randomize the domain parameters"""

    def __init__(self, wrapped_env: Union[SimEnv, EnvWrapper], randomizer: Optional[DomainRandomizer]):
        if not isinstance(inner_env(wrapped_env), SimEnv):
            raise pyrado.TypeErr(given=wrapped_env, expected_type=SimEnv)
        if not isinstance(randomizer, DomainRandomizer) and randomizer is not None:
            raise pyrado.TypeErr(given=randomizer, expected_type=DomainRandomizer)
        super().__init__(wrapped_env)
        self._randomizer = randomizer

    def _ctor_kwargs(self) -> dict:
        return dict(randomizer=self._randomizer)

    @property
    def randomizer(self) -> DomainRandomizer:
        return self._randomizer

    @randomizer.setter
    def randomizer(self, randomizer: DomainRandomizer):
        if not isinstance(randomizer, DomainRandomizer):
            raise pyrado.TypeErr(given=randomizer, expected_type=DomainRandomizer)
        self._randomizer = randomizer


class MetaDomainRandWrapper(DomainRandWrapper):
    """Wraps another `DomainRandWrapper` to adapt its domain distribution parameters (:83-132)."""

    def __init__(self, wrapped_rand_env: DomainRandWrapper, dp_mapping: Mapping[int, Tuple[str, str]]):
        if not typed_env(wrapped_rand_env, DomainRandWrapper):
            raise pyrado.TypeErr(given=wrapped_rand_env, expected_type=DomainRandWrapper)
        super().__init__(wrapped_rand_env, None)
        self.dp_mapping = dp_mapping

    def _ctor_kwargs(self) -> dict:
        return dict(dp_mapping=self.dp_mapping)

    def copy(self, **overrides):
        kw = dict(wrapped_rand_env=overrides.pop("wrapped_env", self._wrapped_env), dp_mapping=self.dp_mapping)
        kw.update(overrides)
        return type(self)(**kw)

    @property
    def randomizer(self) -> DomainRandomizer:
        return self._wrapped_env.randomizer

    @randomizer.setter
    def randomizer(self, dr: DomainRandomizer):
        self._wrapped_env.randomizer = dr

    def adapt_randomizer(self, domain_distr_param_values: np.ndarray):
        vals = np.asarray(domain_distr_param_values)
        if vals.ndim == 2:
            vals = vals.ravel()
        elif vals.ndim != 1:
            raise pyrado.ShapeErr(given=domain_distr_param_values, expected_match=(1,))
        for i, value in enumerate(vals):
            dp_name, ddp_name = self.dp_mapping.get(i)
            self._wrapped_env.randomizer.adapt_one_distr_param(dp_name, ddp_name, value)


class DomainRandWrapperLive(DomainRandWrapper):
    """Randomizes the wrapped env at every reset, thus every rollout is done with different domain parameters."""

    def _reset_chain(self, init_state, domain_param, fusion):
        if domain_param is None:
            sim = inner_env(self)
            if isinstance(sim, SimPyEnv):
                sim.randomize_params(self._randomizer)  # one kernel launch, parameters stay on the device
                extra = [n for n in self._randomizer.names() if n not in sim._param_names]
                if extra:  # wrapper-owned entries (e.g. obs_noise_std): one value for the whole batch
                    self._randomizer.randomize(num_samples=1)
                    one = self._randomizer.get_params(fmt="dict", dtype="numpy")
                    domain_param = {k: one[k] for k in extra}
            else:
                self._randomizer.randomize(num_samples=1)
                domain_param = self._randomizer.get_params(fmt="dict", dtype="numpy")
        return super()._reset_chain(init_state, domain_param, fusion)


class DomainRandWrapperBuffer(DomainRandWrapper):
    """Randomizes the wrapped env from a buffer of domain parameter sets, cycling through it at every reset."""

    def __init__(self, wrapped_env, randomizer: Optional[DomainRandomizer], selection: Optional[str] = "cyclic"):
        if selection not in ["cyclic", "random"]:
            raise pyrado.ValueErr(given=selection, eq_constraint="cyclic or random")
        super().__init__(wrapped_env, randomizer)
        self._ring_idx = None
        self._buffer = None
        self.selection = selection

    def _ctor_kwargs(self) -> dict:
        return dict(randomizer=self._randomizer, selection=self._selection)

    def copy(self, **overrides):
        c = super().copy(**overrides)
        c._buffer, c._ring_idx = self._buffer, self._ring_idx
        return c

    @property
    def ring_idx(self) -> int:
        return self._ring_idx

    @ring_idx.setter
    def ring_idx(self, idx: int):
        if not isinstance(idx, int) or not 0 <= idx < len(self._buffer):
            raise pyrado.ValueErr(given=idx, ge_constraint="0 (int)", l_constraint=len(self._buffer))
        self._ring_idx = idx

    @property
    def selection(self) -> str:
        return self._selection

    @selection.setter
    def selection(self, selection: str):
        if selection not in ["cyclic", "random"]:
            raise pyrado.ValueErr(given=selection, eq_constraint="cyclic or random")
        self._selection = selection

    def fill_buffer(self, num_domains: int):
        """Fill the internal buffer with `num_domains` parameter sets (:202-217)."""
        if self._randomizer is None:
            raise pyrado.TypeErr(msg="The randomizer must not be None to call fill_buffer()!")
        if not isinstance(num_domains, int) or num_domains < 0:
            raise pyrado.ValueErr(given=num_domains, g_constraint="0 (int)")
        self._randomizer.randomize(num_domains)
        self._buffer = self._randomizer.get_params(-1, fmt="list", dtype="numpy")
        self._ring_idx = 0

    @property
    def buffer(self):
        return self._buffer

    @buffer.setter
    def buffer(self, buffer: Union[List[dict], dict]):
        if not isinstance(buffer, (list, dict)):
            raise pyrado.TypeErr(given=buffer, expected_type=[list, dict])
        self._buffer = buffer

    def _next_set(self) -> dict:
        if isinstance(self._buffer, dict):
            return self._buffer
        if isinstance(self._buffer, list):
            dp = self._buffer[self._ring_idx]  # the first selection is index 0
            if self._selection == "cyclic":
                self._ring_idx = (self._ring_idx + 1) % len(self._buffer)
            else:
                self._ring_idx = randint(0, len(self._buffer) - 1)
            return dp
        raise pyrado.TypeErr(given=self._buffer, expected_type=[dict, list])

    def _reset_chain(self, init_state, domain_param, fusion):
        if domain_param is None:
            n = self.num_envs
            if n == 1 or isinstance(self._buffer, dict):
                domain_param = self._next_set()
            else:  # batch: env i takes the i-th next set, the ring advances by num_envs
                sets = [self._next_set() for _ in range(n)]
                domain_param = {k: np.array([float(np.asarray(s[k])) for s in sets]) for k in sets[0]
                                if np.asarray(sets[0][k]).size == 1}
        return super()._reset_chain(init_state, domain_param, fusion)

    # checkpoint hooks (Serializable._get_state/_set_state in the reference, :253-261)
    def state_dict(self) -> dict:
        return dict(buffer=self._buffer, ring_idx=self._ring_idx)

    def load_state_dict(self, state_dict: dict):
        self._buffer, self._ring_idx = state_dict["buffer"], state_dict["ring_idx"]


def remove_all_dr_wrappers(env: Env, verbose: bool = False):
    """Remove every `DomainRandWrapper` from the chain (:264-279)."""
    while any(isinstance(e, DomainRandWrapper) for e in all_envs(env)):
        if verbose:
            print(f"Found domain randomization wrapper of type {type(typed_env(env, DomainRandWrapper)).__name__}. "
                  f"Removing it now.")
        env = remove_env(env, DomainRandWrapper)
    return env
