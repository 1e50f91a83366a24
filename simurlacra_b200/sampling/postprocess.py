"""
Post-processing of sampled batches ON THE DEVICE: discounted returns-to-go and generalised
advantage estimates for every episode of every lane of the time-major record buffer that
``ParallelRolloutSampler.sample_batch(record=True)`` leaves in HBM -- the step after the sampling path in the
reference's actor-critic algorithms (``discounted_values`` / ``gae_returns``, pyrado/sampling/step_sequence.py:1049-1116,
called from algorithms/step_based/gae.py), without copying tens of GB of trajectories to the host first.
"""
from typing import Callable, Optional, Tuple

import torch

from .. import _lib as L
from .. import engine as E


def discounted_returns(batch: dict, gamma: float) -> torch.Tensor:
    """Discounted return-to-go of every step: ``ret [T, N]`` aligned with ``batch['rew']`` (0 on closing / idle rows).
    Equals ``discounted_value(ro, gamma)`` (step_sequence.py:1063-1072) for every complete rollout in the batch."""
    return E.returns(batch["rew"], batch["row_flag"], gamma)[0]


def gae(batch: dict, values: torch.Tensor, gamma: float = 0.99, lamb: float = 0.95) -> Tuple[torch.Tensor, torch.Tensor]:
    """Generalised advantage estimates (``gae_returns``, step_sequence.py:1093-1116) and discounted returns-to-go.

    :param batch: result of ``sample_batch(record=True)`` (device tensors, time-major)
    :param values: value estimates ``[T, N]`` (or ``[T, N, 1]``) of the observation of EVERY row, closing rows included
                   (the reference's ``values`` field has the same ``length + 1`` entries per rollout)
    :return: ``(adv [T, N], ret [T, N])``; episodes cut by the row budget bootstrap from their last value estimate
    """
    if values.dim() == 3:
        values = values[..., 0]
    ret, adv = E.returns(batch["rew"], batch["row_flag"], gamma, values=values, lamb=lamb)
    return adv, ret


def evaluate_values(batch: dict, vfcn: Callable[[torch.Tensor], torch.Tensor], chunk_rows: int = 64) -> torch.Tensor:
    """Apply a torch value function to the observations of all rows of the batch, a few rows at a time (keeps the
    activation memory bounded for 10^8-row batches).  Returns float32 ``[T, N]``."""
    obs = batch["obs"]
    T, N = obs.shape[:2]
    out = torch.empty(T, N, dtype=torch.float32, device=obs.device)
    with torch.no_grad():
        for t0 in range(0, T, chunk_rows):
            o = obs[t0:t0 + chunk_rows].reshape(-1, obs.shape[-1]).to(torch.float32)
            out[t0:t0 + chunk_rows] = vfcn(o).reshape(-1, N)
    return out


def step_mask(batch: dict) -> torch.Tensor:
    """Boolean ``[T, N]``: rows from which a step was taken (the rows that carry actions, rewards, returns, advantages)."""
    return batch["row_flag"] <= L.ROW_DONE
