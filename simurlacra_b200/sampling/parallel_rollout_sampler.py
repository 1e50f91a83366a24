"""
ParallelRolloutSampler (pyrado/sampling/parallel_rollout_sampler.py:182-323) on the CUDA engine.

The reference spawns `num_workers` processes that each run `rollout()` one step at a time and ship pickled
StepSequences back (sampler_pool.py:251-489).  Here the pool is a batch dimension: rollout index `i` of a `sample()`
call is lane `i` of ONE launch of the persistent rollout kernel.  The reference seeds rollout `i` with
`set_seed(seed, sub_seed=sample_count, sub_sub_seed=i)` (parallel_rollout_sampler.py:65-76); the engine keys lane `i`
with Philox `(seed_hash(seed, sample_count), i, episode 0)`.  Hence, like in the reference, the result does not depend
on how the work is split: `num_workers`, the lane chunk size and the number of GPUs (`shard`) leave every rollout
bit-identical (Pyrado/tests/test_sampling.py:570-701).
"""
from itertools import product
from math import ceil
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

import simurlacra_b200 as pyrado

from .. import _lib as L
from .. import engine as E
from ..environments.base import Env
from ..policies.base import Policy
from . import sharding
from .rollout import compile_chain, rollout, sequences_from_records
from .sampler import SamplerBase
from .step_sequence import StepSequence

NO_SEED = object()


class ParallelRolloutSampler(SamplerBase):
    """Class for sampling from many environments in parallel (on the GPU)"""

    def __init__(self, env, policy, num_workers: int = 1, *, min_rollouts: int = None, min_steps: int = None,
                 show_progress_bar: bool = False, seed: int = NO_SEED, shard: Optional[Tuple[int, int]] = None,
                 data_format: str = "numpy", max_lanes: int = 1 << 22):
        """
        :param env: environment (chain) to sample from; its inner `SimPyEnv` supplies the kernel configuration
        :param policy: policy to act in the environment
        :param num_workers: kept for API compatibility; the batch dimension replaces the worker pool
        :param min_rollouts: minimum number of complete rollouts per `sample()`
        :param min_steps: minimum total number of steps per `sample()`
        :param seed: base seed; `NO_SEED` takes the one set with `pyrado.set_seed`
        :param shard: (rank, world_size) -- this process samples only its contiguous block of the rollout indices
                      (one process per GPU); default: the initialised torch.distributed group, else (0, 1)
        :param data_format: 'numpy' (host arrays, reference default) or 'torch' (views of the device buffer)
        :param max_lanes: upper bound of lanes per kernel launch (memory of the record buffer)
        """
        super().__init__(min_rollouts=min_rollouts, min_steps=min_steps)
        if not isinstance(num_workers, int) or num_workers < 1:
            raise pyrado.ValueErr(given=num_workers, ge_constraint="1")
        self.env, self.policy = env, policy
        self.num_workers = num_workers
        self.show_progress_bar = show_progress_bar
        self._seed = pyrado.get_base_seed() if seed is NO_SEED else seed
        self._sample_count = -1
        self._shard = shard
        self.data_format = data_format
        self.max_lanes = max_lanes
        self.last_stats = None  # statistics vector of the last sample() (all-reduced over the shard group)
        self.reinit()

    def reinit(self, env: Optional[Env] = None, policy: Optional[Policy] = None):
        if env is not None:
            self.env = env
        if policy is not None:
            self.policy = policy
        self._cc = compile_chain(self.env)

    # ------------------------------------------------------------------ helpers
    def _shard_rw(self) -> Tuple[int, int]:
        return self._shard if self._shard is not None else sharding.world()

    def _kernel_seed(self) -> int:
        base = self._seed if isinstance(self._seed, int) else 0
        return pyrado.seed_hash(base, self._sample_count, 0)

    def _param_rows(self, sim, dicts: Sequence[dict]) -> torch.Tensor:
        base = sim.base_param_vector()
        rows = np.tile(base, (len(dicts), 1))
        for j, dp in enumerate(dicts):
            for k, v in dp.items():
                if k in sim._param_names:
                    rows[j, sim._param_names.index(k)] = float(np.asarray(v))
        return torch.as_tensor(rows, dtype=sim.dtype).to(sim.device).contiguous()

    def _init_rows(self, sim, states: Sequence[np.ndarray]) -> torch.Tensor:
        arr = np.stack([np.asarray(s, dtype=np.float64).reshape(-1) for s in states])
        if arr.shape[1] not in (sim._dims["ds"], sim._dims["dinit"]):
            raise pyrado.ShapeErr(given=arr[0], expected_match=sim.state_space)
        return torch.as_tensor(arr, dtype=sim.dtype).to(sim.device).contiguous()

    def _launch(self, lo: int, hi: int, desc, init=None, params=None, record=True, stats=None) -> dict:
        """Rollout indices [lo, hi) as one kernel launch (lane j = global rollout index lo + j)."""
        cc = self._cc
        sim = cc.sim
        distr = None
        if params is None:
            params = sim._params if (sim._params is not None and sim._params.shape[0] == hi - lo) else None
            if cc.live_randomizer is not None:
                distr = cc.live_randomizer.table(sim._param_names)
        cfg = sim._make_cfg(cc.fusion)
        res = E.rollout(cfg, desc, hi - lo, int(sim.max_steps) + 1, device=sim.device, params=params, distr=distr,
                        init=init, seed=self._kernel_seed(), env_offset=lo, episode=0, stop_on_done=True,
                        auto_reset=False, record=record, want_stats=True)
        if stats is not None:
            stats[: L.STAT_MIN_RET] += res["stats"][: L.STAT_MIN_RET]
            stats[L.STAT_MIN_RET] = torch.minimum(stats[L.STAT_MIN_RET], res["stats"][L.STAT_MIN_RET])
            stats[L.STAT_MAX_RET] = torch.maximum(stats[L.STAT_MAX_RET], res["stats"][L.STAT_MAX_RET])
        return res

    def _check_status(self, res):
        st = int(res["status"].max()) if res["status"].numel() else 0
        if st & (L.ST_NAN_OBS | L.ST_NAN_STATE):
            raise pyrado.ValueErr(msg="At least one observation value is NaN!")
        if st & L.ST_NAN_ACT:
            raise pyrado.ValueErr(msg="At least one action value is NaN!")

    # ------------------------------------------------------------------ sample
    def sample(self, init_states: Optional[List[np.ndarray]] = None, domain_params: Optional[List[dict]] = None,
               eval: bool = False) -> List[StepSequence]:
        """
        Sample rollouts until the minimum counts are met (parallel_rollout_sampler.py:248-323).

        :param init_states: initial states forwarded to `env.reset()`; repeated until `min_rollouts` is reached
        :param domain_params: domain parameter dicts forwarded to `env.reset()`; given together with `init_states`
                              every combination is sampled
        :param eval: pass `False` if the rollout is executed during training, else `True`
        :return: list of sampled rollouts (this rank's block when sharded)
        """
        self._sample_count += 1
        cc = self._cc = compile_chain(self.env)
        desc = None
        if cc is not None and isinstance(self.policy, Policy) and np.isfinite(self.env.max_steps):
            desc = self.policy.engine_desc(cc.sim.device)  # fresh copy of the weights: the policy update
        if desc is None:
            return self._sample_generic(init_states, domain_params, eval)
        self.policy.eval() if eval else self.policy.train()
        sim = cc.sim
        rank, ws = self._shard_rw()
        stats = E.stats_init(sim.device)

        if self.min_steps is None:
            # ---- a fixed number of rollouts: one index per (repeated) reset argument
            init_all = params_all = None
            if init_states is None and domain_params is None:
                n_total = self.min_rollouts
            else:
                if init_states is not None and domain_params is not None:
                    combos = list(product(init_states, domain_params))
                    rep = ceil(self.min_rollouts / len(combos)) * combos
                    init_list, dp_list = [c[0] for c in rep], [c[1] for c in rep]
                elif init_states is not None:
                    init_list, dp_list = ceil(self.min_rollouts / len(init_states)) * list(init_states), None
                else:
                    init_list, dp_list = None, ceil(self.min_rollouts / len(domain_params)) * list(domain_params)
                n_total = len(init_list if init_list is not None else dp_list)
                init_all = self._init_rows(sim, init_list) if init_list is not None else None
                params_all = self._param_rows(sim, dp_list) if dp_list is not None else None
            lo, hi = sharding.shard_range(n_total, rank, ws)
            ros = []
            for a in range(lo, hi, self.max_lanes):
                b = min(a + self.max_lanes, hi)
                res = self._launch(a, b, desc, None if init_all is None else init_all[a:b].contiguous(),
                                   None if params_all is None else params_all[a:b].contiguous(), stats=stats)
                self._check_status(res)
                ros.extend(sequences_from_records(self.env, sim, res, sim.dt, self.data_format, obs_layout=cc.obs_layout))
        else:
            if init_states is not None:
                raise NotImplementedError  # like the reference (parallel_rollout_sampler.py:319-320)
            # ---- a minimum number of steps: rollout indices 0, 1, 2, ... in order until both limits hold; the
            # result is cut after the first index that satisfies them, so it does not depend on the chunking
            horizon = int(sim.max_steps)
            # lanes per launch: episodes rarely use the whole horizon, and a launch of a few dozen lanes costs the same ~2 ms as
            # one of a few hundred (it is latency-bound), so the first launch over-provisions four-fold -- the result is cut
            # after the first index that satisfies the limits either way
            chunk = min(self.max_lanes, max(self.min_rollouts or 1, 4 * ceil(self.min_steps / horizon)))
            lengths, results, done_n = [], [], 0
            n_keep = None
            while n_keep is None:
                res = self._launch(done_n, done_n + chunk, desc, record=True)
                self._check_status(res)
                results.append((done_n, res))
                lengths.append(res["length"].cpu().numpy())
                done_n += chunk
                csum = np.cumsum(np.concatenate(lengths))
                ok = np.flatnonzero(csum >= self.min_steps)
                if len(ok) and (self.min_rollouts is None or done_n >= self.min_rollouts):
                    n_keep = max(int(ok[0]) + 1, self.min_rollouts or 0)
            lo, hi = sharding.shard_range(n_keep, rank, ws)
            ros = []
            st = np.array([0, 0, 0, 0, 0, 0, 1e300, -1e300], dtype=np.float64)
            for start, res in results:
                ids = np.arange(max(lo, start), min(hi, start + chunk)) - start
                if len(ids):
                    ros.extend(sequences_from_records(self.env, sim, res, sim.dt, self.data_format, env_ids=ids,
                                                      obs_layout=cc.obs_layout))
                    # statistics of the kept rollouts only, from the kernel's per-lane results (no per-rollout pass)
                    r = res["ret"].double().cpu().numpy()[ids]
                    ln = res["length"].cpu().numpy()[ids].astype(np.float64)
                    flag = res["row_flag"].cpu().numpy()
                    done = flag[ln.astype(np.int64) - 1, ids] == L.ROW_DONE
                    st[L.STAT_EPISODES] += len(ids)
                    st[L.STAT_SUM_RET] += r.sum()
                    st[L.STAT_SUM_RET2] += (r * r).sum()
                    st[L.STAT_SUM_LEN] += ln.sum()
                    st[L.STAT_STEPS] += ln.sum()
                    st[L.STAT_FAILED] += float((done & (ln < horizon)).sum())
                    st[L.STAT_MIN_RET] = min(st[L.STAT_MIN_RET], r.min())
                    st[L.STAT_MAX_RET] = max(st[L.STAT_MAX_RET], r.max())
            stats = torch.as_tensor(st, device=sim.device)
        self.last_stats = sharding.allreduce_stats(stats) if self._shard is None else stats
        return ros

    def statistics(self) -> dict:
        """avg / std / min / max return and lengths of the last `sample()` over ALL ranks (engine extension)."""
        return sharding.summarize(self.last_stats)

    def _sample_generic(self, init_states, domain_params, eval) -> List[StepSequence]:
        """Chains / policies the kernel cannot fuse: the reference's per-index loop on the step kernels."""
        def one(i, **reset_kwargs):
            ro = rollout(self.env, self.policy, eval=eval, seed=self._seed if isinstance(self._seed, int) else 0,
                         sub_seed=self._sample_count, sub_sub_seed=i, reset_kwargs=reset_kwargs or None)
            return ro if isinstance(ro, list) else [ro]

        ros = []
        if self.min_steps is None:
            if init_states is None and domain_params is None:
                args = [dict() for _ in range(self.min_rollouts)]
            elif domain_params is None:
                args = [dict(init_state=s) for s in ceil(self.min_rollouts / len(init_states)) * list(init_states)]
            elif init_states is None:
                args = [dict(domain_param=d) for d in ceil(self.min_rollouts / len(domain_params)) * list(domain_params)]
            else:
                combos = list(product(init_states, domain_params))
                args = [dict(init_state=s, domain_param=d) for s, d in ceil(self.min_rollouts / len(combos)) * combos]
            for i, kw in enumerate(args):
                ros.extend(one(i, **kw))
        else:
            if init_states is not None:
                raise NotImplementedError
            i = steps = 0
            while steps < self.min_steps or (self.min_rollouts is not None and len(ros) < self.min_rollouts):
                new = one(i)
                ros.extend(new)
                steps += sum(len(r) for r in new)
                i += 1
        return ros

    # ------------------------------------------------------------------ engine extension: no host round trip
    def sample_batch(self, num_envs: int, *, auto_reset: bool = True, record: bool = False, eval: bool = True,
                     traj: Optional[torch.Tensor] = None, row_flag: Optional[torch.Tensor] = None,
                     gather: Optional["sharding.RecordGather"] = None) -> dict:
        """
        One launch for `num_envs` lanes on this rank, results left ON THE DEVICE: returns / lengths / parameters of
        each lane's first episode, the all-reduced statistics vector and (with `record`) the time-major record buffer.
        With `auto_reset` every lane keeps starting new episodes until the horizon is used up (`min_steps`-style
        sampling: exactly `num_envs * max_steps` steps).  This is the call the benchmark times.
        With `gather` (a `sharding.RecordGather` over num_envs * world_size rollouts) the records of all ranks are written
        straight into the learner rank's buffer by the rollout kernels (NVLink peer stores); call `gather.finish()` after.
        """
        self._sample_count += 1
        cc = self._cc = compile_chain(self.env)
        if cc is None:
            raise L.EngineError("sample_batch needs an env chain that fuses into the rollout kernel")
        desc = self.policy.engine_desc(cc.sim.device)
        if desc is None:
            raise L.EngineError(f"{type(self.policy).__name__} has no fused CUDA forward pass")
        sim = cc.sim
        rank, ws = self._shard_rw()
        distr = cc.live_randomizer.table(sim._param_names) if cc.live_randomizer is not None else None
        res = E.rollout(sim._make_cfg(cc.fusion), desc, num_envs, int(sim.max_steps) + 1, device=sim.device,
                        params=None if distr is not None else (sim._params if sim._params is not None and
                                                               sim._params.shape[0] == num_envs else None),
                        distr=distr, seed=self._kernel_seed(), env_offset=rank * num_envs, episode=0,
                        stop_on_done=True, auto_reset=auto_reset, record=record or gather is not None,
                        **(gather.target() if gather is not None else dict(traj=traj, row_flag=row_flag)))
        self.last_stats = sharding.allreduce_stats(res["stats"]) if self._shard is None else res["stats"]
        return res


    def read_results(self, res: dict):
        """Returns, lengths and the statistics vector of a `sample_batch` result on the HOST, through pinned staging buffers
        the sampler keeps (one asynchronous copy each and one synchronisation; a fresh pageable `.cpu()` per tensor costs
        ~4 ms for 2^20 lanes, this ~0.3 ms).  The arrays are views of the staging buffers: valid until the next call."""
        n = res["ret"].shape[0]
        st = getattr(self, "_host_stage", None)
        if st is None or st[0].shape[0] != n or st[0].dtype != res["ret"].dtype:
            st = self._host_stage = (torch.empty(n, dtype=res["ret"].dtype).pin_memory(),
                                     torch.empty(n, dtype=torch.int32).pin_memory(),
                                     torch.empty(L.N_STATS, dtype=torch.float64).pin_memory())
        st[0].copy_(res["ret"], non_blocking=True)
        st[1].copy_(res["length"], non_blocking=True)
        st[2].copy_(self.last_stats, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return st[0].numpy(), st[1].numpy(), st[2].numpy()


class ParallelRolloutSamplerTensor(SamplerBase):
    """The fork's batched sampler for differentiable environments (parallel_rollout_sampler.py:326-500): `sample()` calls
    `rollout_tensor_full(env, policy, bsz=32, max_steps=min_steps)` until `min_steps` steps (and `min_rollouts` rollouts)
    have been collected; every StepSequence carries the per-step Jacobians in `env_infos`.  Here the lanes are a batch on
    the GPU and `bsz` is a constructor argument (the reference hard-codes 32)."""

    def __init__(self, env, policy, num_workers: int = 1, *, min_rollouts: int = None, min_steps: int = None,
                 show_progress_bar: bool = False, seed: int = NO_SEED, bsz: int = 32):
        super().__init__(min_rollouts=min_rollouts, min_steps=min_steps)
        if not isinstance(num_workers, int) or num_workers < 1:
            raise pyrado.ValueErr(given=num_workers, ge_constraint="1")
        self.env, self.policy, self.num_workers, self.bsz = env, policy, num_workers, int(bsz)
        self.show_progress_bar = show_progress_bar
        self._seed = pyrado.get_base_seed() if seed is NO_SEED else seed
        self._sample_count = -1

    def reinit(self, env: Optional[Env] = None, policy: Optional[Policy] = None):
        if env is not None:
            self.env = env
        if policy is not None:
            self.policy = policy

    def sample(self, init_states: Optional[List[np.ndarray]] = None, domain_params: Optional[List[dict]] = None,
               eval: bool = False) -> List[StepSequence]:
        from .rollout_tensor import rollout_tensor_full

        self._sample_count += 1
        if init_states is not None or domain_params is not None or self.min_steps is None:
            raise NotImplementedError  # like the reference: only the min_steps mode without reset arguments exists
        result, counter = [], 0
        while counter < self.min_steps or (self.min_rollouts is not None and len(result) < self.min_rollouts):
            res = rollout_tensor_full(self.env, self.policy, eval=eval, seed=self._seed if isinstance(self._seed, int) else 0,
                                      sub_seed=self._sample_count, sub_sub_seed=len(result) + 1, bsz=self.bsz,
                                      max_steps=self.min_steps)
            result += res
            counter += sum(len(r) for r in res)
        return result
