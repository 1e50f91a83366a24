"""
Multi-GPU plumbing of the sampler: environments (rollout indices) are the data-parallel axis and shard
with NO exchange during the horizon -- the Philox key uses the GLOBAL rollout index, so any split over ranks yields
the same rollouts (the analogue of the reference's "same result for 1, 2 or 4 workers",
Pyrado/tests/test_sampling.py:594-646).  Collectives appear only after a `sample()`:
  - one all-gather of the 64-byte statistics vector, combined locally into sums / min / max (what the learner logs,
    algorithms/step_based/actor_critic.py:115-121)
  - an optional all-gather of per-rollout returns / lengths (ragged over ranks -> padded)
Works with NCCL on GPUs and with gloo on CPU tensors (tests).
"""
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from .. import _lib as L


def world(group=None) -> Tuple[int, int]:
    """(rank, world_size); (0, 1) without an initialised process group."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_range(n_total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of the global index range [0, n_total) owned by `rank`; sizes differ by at most 1."""
    base, rem = divmod(n_total, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_stats(stats: torch.Tensor, group=None) -> torch.Tensor:
    """In-place reduction of a [SRE_N_STATS] float64 statistics vector over the ranks: sums for the first six entries, min /
    max after.  ONE collective: the 64-byte vectors are all-gathered and combined locally (sum, min and max would need
    three all-reduces; at this size a collective is pure launch latency)."""
    ws = world(group)[1]
    if ws == 1:
        return stats
    flat = torch.empty(ws * stats.shape[0], dtype=stats.dtype, device=stats.device)  # flat: gloo wants the concatenated form
    dist.all_gather_into_tensor(flat, stats.contiguous(), group=group)
    allv = flat.view(ws, stats.shape[0])
    stats[: L.STAT_MIN_RET] = allv[:, : L.STAT_MIN_RET].sum(dim=0)
    stats[L.STAT_MIN_RET] = allv[:, L.STAT_MIN_RET].min()
    stats[L.STAT_MAX_RET] = allv[:, L.STAT_MAX_RET].max()
    return stats


def summarize(stats: torch.Tensor) -> dict:
    """avg / std / min / max return, avg rollout length, counts (what ActorCritic.step logs)."""
    s = stats.detach().cpu().tolist()
    n = max(s[L.STAT_EPISODES], 1.0)
    mean = s[L.STAT_SUM_RET] / n
    var = max(s[L.STAT_SUM_RET2] / n - mean * mean, 0.0)
    return dict(num_rollouts=int(s[L.STAT_EPISODES]), num_steps=int(s[L.STAT_STEPS]), num_failed=int(s[L.STAT_FAILED]),
                avg_return=mean, std_return=var ** 0.5, min_return=s[L.STAT_MIN_RET], max_return=s[L.STAT_MAX_RET],
                avg_rollout_len=s[L.STAT_SUM_LEN] / n)


def allgather_ragged(x: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """Concatenate the per-rank blocks of a sharded [n_local, ...] tensor into [n_total, ...] on every rank, in global
    index order (blocks as cut by `shard_range`)."""
    rank, ws = world(group)
    if ws == 1:
        return x
    n_max = -(-n_total // ws)
    pad = torch.zeros(n_max, *x.shape[1:], dtype=x.dtype, device=x.device)
    pad[: x.shape[0]] = x
    parts = [torch.empty_like(pad) for _ in range(ws)]
    dist.all_gather(parts, pad, group=group)
    out = []
    for r, p in enumerate(parts):
        lo, hi = shard_range(n_total, r, ws)
        out.append(p[: hi - lo])
    return torch.cat(out)


def broadcast_policy(policy: torch.nn.Module, src: int = 0, group=None):
    """Replicate the policy parameters from `src` (analogue of `_ps_update_policy`, parallel_rollout_sampler.py:60-62)."""
    if world(group)[1] == 1:
        return
    for p in policy.parameters():
        dist.broadcast(p.data, src=src, group=group)
