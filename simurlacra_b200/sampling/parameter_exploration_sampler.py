"""
ParameterExplorationSampler (pyrado/sampling/parameter_exploration_sampler.py:62-345): evaluate MANY policy-parameter
sets, each on the same `num_domains` domain-parameter sets x `num_init_states_per_domain` initial states (common
random numbers across the population), as the episodic algorithms (CEM, NES, PoWER, REPS, hill climbing) do.

The reference pickles the env / policy into a SamplerPool and writes every parameter vector into the worker's policy
before each rollout.  Here the whole population is ONE launch of the rollout kernel: lane n = (parameter set
n // R, repetition n % R) with R = num_domains * num_init_states_per_domain; the parameter sets live in one
[pop, num_param] float32 matrix on the device and each lane reads its own set (`sre_policy.group_size`).
"""
from typing import List, NamedTuple, Optional, Sequence, Tuple

import numpy as np
import torch

import simurlacra_b200 as pyrado

from .. import _lib as L
from .. import engine as E
from ..environment_wrappers.domain_randomization import (DomainRandWrapper, DomainRandWrapperBuffer,
                                                         remove_all_dr_wrappers)
from ..environment_wrappers.utils import inner_env, typed_env
from ..environments.sim_base import SimEnv
from ..policies.base import Policy
from . import sharding
from .rollout import compile_chain, rollout, sequences_from_records
from .step_sequence import StepSequence

NO_SEED = object()


class ParameterSample(NamedTuple):
    """Stores policy parameters and associated rollouts."""

    params: torch.Tensor
    rollouts: List[StepSequence]

    @property
    def mean_undiscounted_return(self) -> float:
        return float(np.mean([r.undiscounted_return() for r in self.rollouts]))

    @property
    def num_rollouts(self) -> int:
        return len(self.rollouts)


class ParameterSamplingResult(Sequence):
    """Result of a parameter exploration sampling run: a list of ParameterSamples that also hands out the parameters
    and the mean returns as combined arrays (:79-124)."""

    def __init__(self, samples: Sequence[ParameterSample], mean_returns: Optional[np.ndarray] = None):
        self._samples = samples
        self._mean_returns = mean_returns  # engine extension: already reduced on the device

    def __getitem__(self, idx):
        res = self._samples[idx]
        if not isinstance(res, ParameterSample):
            mr = None if self._mean_returns is None else self._mean_returns[idx]
            return ParameterSamplingResult(res, mr)
        return res

    def __len__(self):
        return len(self._samples)

    @property
    def parameters(self) -> torch.Tensor:
        """All policy parameters as an N x P matrix"""
        return torch.stack([s.params for s in self._samples])

    @property
    def mean_returns(self) -> np.ndarray:
        """The mean return of every parameter sample"""
        if self._mean_returns is not None:
            return np.asarray(self._mean_returns)
        return np.array([s.mean_undiscounted_return for s in self._samples])

    @property
    def rollouts(self) -> list:
        return [s.rollouts for s in self._samples]

    @property
    def num_rollouts(self) -> int:
        return int(sum(s.num_rollouts for s in self._samples))


class ParameterExplorationSampler:
    """Parallel sampler for parameter exploration"""

    def __init__(self, env, policy: Policy, num_init_states_per_domain: int, num_domains: int, num_workers: int = 1,
                 seed: Optional[int] = NO_SEED, shard: Optional[Tuple[int, int]] = None):
        """
        :param env: environment to sample from; a DomainRandWrapper around it provides the domains
        :param policy: policy whose parameters are replaced by the sampled sets
        :param num_init_states_per_domain: number of rollouts per domain
        :param num_domains: number of domain-parameter sets per policy-parameter set
        :param num_workers: kept for the reference's signature (the batch dimension replaces the process pool)
        :param seed: base seed; rollout `i` of the `k`-th `sample()` call uses `(seed, k, i)` like the reference
        :param shard: (rank, world_size) -- engine extension, one process per GPU: this process rolls out only its
                      contiguous block of the parameter sets (no exchange during the horizon; domains, initial states
                      and noise streams are keyed on the seed and the global lane index, so every split gives the same
                      rollouts) and the mean returns are all-gathered afterwards.  Default: the initialised
                      torch.distributed group, else (0, 1).  Needs an integer seed shared by all ranks.
        """
        if not isinstance(num_init_states_per_domain, int):
            raise pyrado.TypeErr(given=num_init_states_per_domain, expected_type=int)
        if num_init_states_per_domain < 1:
            raise pyrado.ValueErr(given=num_init_states_per_domain, ge_constraint="1")
        if not isinstance(num_domains, int):
            raise pyrado.TypeErr(given=num_domains, expected_type=int)
        if num_domains < 1:
            raise pyrado.ValueErr(given=num_domains, ge_constraint="1")
        self._dr_wrapper = typed_env(env, DomainRandWrapper)
        if self._dr_wrapper is not None:
            assert isinstance(inner_env(env), SimEnv)
            env = remove_all_dr_wrappers(env)
        self.env, self.policy = env, policy
        self.num_init_states_per_domain = num_init_states_per_domain
        self.num_domains = num_domains
        self.num_workers = num_workers
        self._seed = pyrado.get_base_seed() if seed is NO_SEED else seed
        self._sample_count = -1
        self._shard = shard

    @property
    def num_rollouts_per_param(self) -> int:
        return self.num_init_states_per_domain * self.num_domains

    def reinit(self, env=None, policy: Optional[Policy] = None):
        if env is not None:
            self.env = env
        if policy is not None:
            self.policy = policy

    # ------------------------------------------------------------------ the shared domains and init states
    def _sample_domain_params(self) -> list:
        """One list entry per domain (:218-232): None without a DR wrapper, the buffer's sets, or fresh draws."""
        if self._dr_wrapper is None:
            return [None] * self.num_domains
        if isinstance(self._dr_wrapper, DomainRandWrapperBuffer) and self._dr_wrapper.buffer is not None:
            idcs = np.random.randint(0, len(self._dr_wrapper.buffer), size=self.num_domains)
            return [self._dr_wrapper.buffer[i] for i in idcs]
        key = pyrado.seed_hash(self._seed if isinstance(self._seed, int) else 0, self._sample_count, 1)
        self._dr_wrapper.randomizer.randomize(self.num_domains, seed=key, episode=self._sample_count)
        return self._dr_wrapper.randomizer.get_params(-1, fmt="list", dtype="numpy")

    def _sample_one_init_state(self, domain_param: Optional[dict]) -> Optional[np.ndarray]:
        """The init space may depend on the domain parameters: set them, then sample (:234-249)."""
        self.env.reset(domain_param=domain_param)
        ispace = getattr(self.env, "init_space", None)
        return None if ispace is None else ispace.sample_uniform()

    # ------------------------------------------------------------------ sample
    def sample(self, param_sets: torch.Tensor, init_states: Optional[List[np.ndarray]] = None,
               return_rollouts: bool = True) -> ParameterSamplingResult:
        """
        Sample rollouts for a given set of parameters (:267-345).

        :param param_sets: policy parameter sets, `[pop_size, num_param]`
        :param init_states: one initial state per (domain, repetition), or `None` to sample them
        :param return_rollouts: engine extension -- `False` skips cutting the `StepSequence`s out of the record buffer
                                (the result then only carries parameters and mean returns, all an episodic algorithm's
                                update needs)
        """
        if init_states is not None and not isinstance(init_states, list):
            raise pyrado.TypeErr(given=init_states, expected_type=list)
        param_sets = torch.as_tensor(param_sets)
        if param_sets.dim() != 2 or param_sets.shape[1] != self.policy.num_param:
            raise pyrado.ShapeErr(msg=f"param_sets must be [pop_size, {self.policy.num_param}], got {tuple(param_sets.shape)}")
        self._sample_count += 1
        domain_params = self._sample_domain_params()
        R, P = self.num_rollouts_per_param, param_sets.shape[0]
        if init_states is None:
            # the spaces draw from numpy's global generator: key it on (seed, call count) for the draw and put the
            # caller's stream back, so that a seeded sampler is reproducible without side effects
            keep = np.random.get_state()
            if isinstance(self._seed, int):
                np.random.seed(pyrado.seed_hash(self._seed, self._sample_count, 2) % (1 << 32))
            try:
                init_states = [self._sample_one_init_state(dp) for dp in domain_params
                               for _ in range(self.num_init_states_per_domain)]
            finally:
                if isinstance(self._seed, int):
                    np.random.set_state(keep)
        elif not len(init_states) == R:
            raise pyrado.ShapeErr(msg=f"{len(init_states)} initial states given, {R} needed")
        domain_params = domain_params * self.num_init_states_per_domain

        cc = compile_chain(self.env)
        sim = cc.sim if cc is not None else None
        desc = None
        rank, ws = self._shard if self._shard is not None else sharding.world()
        lo, hi = sharding.shard_range(P, rank, ws)  # this rank's block of the population
        if cc is not None and np.isfinite(self.env.max_steps) and all(s is not None for s in init_states) and hi > lo:
            sets_dev = param_sets[lo:hi].detach().to(device=sim.device, dtype=torch.float32).contiguous()
            desc = self.policy.engine_desc(sim.device, param_sets=sets_dev, group_size=R)
        if desc is None and ws > 1 and cc is not None:
            raise pyrado.ValueErr(msg="Sharded parameter exploration needs a policy and a chain the kernels implement "
                                      "and at least one parameter set per rank!")
        if desc is None:
            return self._sample_generic(param_sets, domain_params, init_states)

        # lane n = (set n // R, repetition n % R): domain parameters and init states repeat with period R
        base = sim.base_param_vector()
        prm = np.tile(base, (R, 1))
        for r, dp in enumerate(domain_params):
            if dp is not None:
                for k, v in dp.items():
                    if k in sim._param_names:
                        prm[r, sim._param_names.index(k)] = float(np.asarray(v).reshape(-1)[0])
        ini = np.stack([np.asarray(s, dtype=np.float64).reshape(-1) for s in init_states])
        Pl = hi - lo
        to_dev = lambda a: torch.as_tensor(a, dtype=sim._dtype).to(sim.device).repeat(Pl, 1).contiguous()
        seed = pyrado.seed_hash(self._seed if isinstance(self._seed, int) else 0, self._sample_count, 0)
        res = E.rollout(sim._make_cfg(cc.fusion), desc, Pl * R, int(sim.max_steps) + 1, device=sim.device,
                        params=to_dev(prm), init=to_dev(ini), seed=seed, env_offset=lo * R, episode=0, stop_on_done=True,
                        auto_reset=False, record=return_rollouts, want_stats=False)
        st = int(res["status"].max())
        if st & (L.ST_NAN_OBS | L.ST_NAN_STATE):
            raise pyrado.ValueErr(msg="At least one observation value is NaN!")
        if st & L.ST_NAN_ACT:
            raise pyrado.ValueErr(msg="At least one action value is NaN!")
        mean_ret = res["ret"].double().reshape(Pl, R).mean(dim=1)
        if ws > 1 and self._shard is None:  # the only exchange: pop_size doubles, after the horizon
            mean_ret = sharding.allgather_ragged(mean_ret, P)
        elif ws > 1:  # explicit shard without a process group: the caller merges; foreign sets stay NaN
            full = torch.full((P,), float("nan"), dtype=torch.float64, device=mean_ret.device)
            full[lo:hi] = mean_ret
            mean_ret = full
        ros = sequences_from_records(self.env, sim, res, float(sim.dt), obs_layout=cc.obs_layout) if return_rollouts else []
        samples = [ParameterSample(params=param_sets[p],
                                   rollouts=ros[(p - lo) * R:(p - lo + 1) * R] if return_rollouts and lo <= p < hi else [])
                   for p in range(P)]
        return ParameterSamplingResult(samples, mean_ret.cpu().numpy())

    def _sample_generic(self, param_sets, domain_params, init_states) -> ParameterSamplingResult:
        """Chains / policies the kernel cannot fuse: the reference's loop, one rollout() per (set, domain, init)."""
        samples, num = [], 0
        seed = self._seed if isinstance(self._seed, int) else 0
        for p in param_sets:
            self.policy.param_values = p.to(self.policy.param_values)
            ros = []
            for dp, s0 in zip(domain_params, init_states):
                ros.append(rollout(self.env, self.policy, reset_kwargs=dict(init_state=s0, domain_param=dp), seed=seed,
                                   sub_seed=self._sample_count, sub_sub_seed=num))
                num += 1
            samples.append(ParameterSample(params=p, rollouts=ros))
        return ParameterSamplingResult(samples)
