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


# ---------------------------------------------------------------------------------------------------------------------
# Rollouts for the learner: the record shards of all ranks in one buffer on one rank
# ---------------------------------------------------------------------------------------------------------------------
def gather_records(traj: torch.Tensor, row_flag: torch.Tensor, n_total: int, dst: int = 0, group=None):
    """Collective baseline (NCCL / gloo): every rank's time-major record shard [T, n_local, W] and flags [T, n_local] end
    up on `dst` as [T, n_total, W] / [T, n_total] in global rollout order (blocks as cut by `shard_range`).  Returns
    (traj, row_flag) on `dst`, (None, None) elsewhere.  The transfer FOLLOWS the simulation; `RecordGather` below makes
    the rollout kernels write into the learner's memory directly instead."""
    rank, ws = world(group)
    if ws == 1:
        return traj, row_flag
    T, _, W = traj.shape
    n_max = -(-n_total // ws)
    pad_t = torch.zeros(T, n_max, W, dtype=traj.dtype, device=traj.device)
    pad_f = torch.zeros(T, n_max, dtype=row_flag.dtype, device=row_flag.device)
    pad_t[:, : traj.shape[1]] = traj
    pad_f[:, : row_flag.shape[1]] = row_flag
    parts_t = [torch.empty_like(pad_t) for _ in range(ws)] if rank == dst else None
    parts_f = [torch.empty_like(pad_f) for _ in range(ws)] if rank == dst else None
    dist.gather(pad_t, parts_t, dst=dst, group=group)
    dist.gather(pad_f, parts_f, dst=dst, group=group)
    if rank != dst:
        return None, None
    out_t = torch.empty(T, n_total, W, dtype=traj.dtype, device=traj.device)
    out_f = torch.empty(T, n_total, dtype=row_flag.dtype, device=row_flag.device)
    for r in range(ws):
        lo, hi = shard_range(n_total, r, ws)
        out_t[:, lo:hi] = parts_t[r][:, : hi - lo]
        out_f[:, lo:hi] = parts_f[r][:, : hi - lo]
    return out_t, out_f


class RecordGather:
    """The learner's record buffer as the DIRECT target of every rank's rollout launch ("gather rollouts for the
    learner"): `traj` [T, n_total, W] and `row_flag` [T, n_total] live on rank `dst`, every other rank maps them through
    torch's symmetric memory (CUDA peer mappings over NVLink / NVSwitch) and passes the mapped pointers to `sre_rollout`
    with its column offset (`sre_rollout_io.traj_env_offset`).  The kernel's record tiles -- 128-bit streaming stores or
    bulk (TMA) stores of a warp's 32 records -- then travel over NVLink while the simulation runs: one kernel does the
    compute and the transfer, no collective follows.  `finish()` is the only synchronisation (a device-side barrier on
    the launching stream after the kernel).

    Every rank allocates the symmetric buffer (the allocation is collective and symmetric), so the size has to fit one
    GPU: T * n_total * (W * 4 + 1) bytes."""

    def __init__(self, n_rows: int, n_total: int, width: int, dtype=torch.float32, device=None, dst: int = 0, group=None):
        self.rank, self.ws = world(group)
        self.dst, self.n_total, self.group = dst, n_total, group
        self.lo, self.hi = shard_range(n_total, self.rank, self.ws)
        self._hdl_t = self._hdl_f = None
        if self.ws == 1:
            self.traj = torch.empty(n_rows, n_total, width, dtype=dtype, device=device)
            self.row_flag = torch.empty(n_rows, n_total, dtype=torch.uint8, device=device)
            return
        import torch.distributed._symmetric_memory as symm

        g = group if group is not None else dist.group.WORLD
        self._local_t = symm.empty((n_rows, n_total, width), dtype=dtype, device=device)
        self._local_f = symm.empty((n_rows, n_total), dtype=torch.uint8, device=device)
        self._hdl_t = symm.rendezvous(self._local_t, group=g)
        self._hdl_f = symm.rendezvous(self._local_f, group=g)
        self.traj = self._hdl_t.get_buffer(dst, (n_rows, n_total, width), dtype)
        self.row_flag = self._hdl_f.get_buffer(dst, (n_rows, n_total), torch.uint8)

    def target(self) -> dict:
        """Keyword arguments for engine.rollout / ParallelRolloutSampler.sample_batch of THIS rank."""
        return dict(traj=self.traj, row_flag=self.row_flag, traj_env_offset=self.lo)

    def finish(self):
        """After the rollout launch(es): every rank's records have arrived on `dst` once this returns there."""
        if self._hdl_t is not None:
            self._hdl_t.barrier(channel=0)
        return (self.traj, self.row_flag) if self.rank == self.dst else (None, None)
