"""
A PPO-style training loop that never leaves the GPU: sampling with exploration in the rollout kernel
(``ParallelRolloutSampler.sample_batch`` with a ``NormalActNoiseExplStrat``), value estimates with a torch value
network on the record buffer, returns-to-go and GAE with ``sre_returns`` (``sampling.postprocess``), a clipped
policy-gradient update with torch autograd on the device tensors.  The reference runs the same steps through
``ParallelRolloutSampler.sample`` -> ``StepSequence`` lists -> ``gae_returns`` on the host
(algorithms/step_based/{ppo,gae}.py).

    python examples/ppo_on_device.py [--iters 20] [--envs 4096]
"""
import argparse
import os.path as osp
import sys

import numpy as np
import torch

sys.path.insert(0, osp.dirname(osp.dirname(osp.abspath(__file__))))

from simurlacra_b200.environment_wrappers import ActNormWrapper  # noqa: E402
from simurlacra_b200.environments.pysim.one_mass_oscillator import OneMassOscillatorSim  # noqa: E402
from simurlacra_b200.exploration import NormalActNoiseExplStrat  # noqa: E402
from simurlacra_b200.policies import FNNPolicy  # noqa: E402
from simurlacra_b200.sampling import ParallelRolloutSampler, postprocess  # noqa: E402


def train(iters: int = 20, n_envs: int = 4096, max_steps: int = 100, seed: int = 0, verbose: bool = True):
    torch.manual_seed(seed)
    dev = torch.device("cuda", 0)
    env = ActNormWrapper(OneMassOscillatorSim(dt=0.02, max_steps=max_steps, task_args=dict(state_des=np.array([0.5, 0.0])),
                                              device=dev))
    policy = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh, use_cuda=True)
    expl = NormalActNoiseExplStrat(policy, std_init=0.3)
    vfcn = torch.nn.Sequential(torch.nn.Linear(2, 32), torch.nn.Tanh(), torch.nn.Linear(32, 1)).to(dev)
    sampler = ParallelRolloutSampler(env, expl, num_workers=1, min_rollouts=1, seed=seed)
    opt_pi = torch.optim.Adam(expl.parameters(), lr=3e-3)
    opt_v = torch.optim.Adam(vfcn.parameters(), lr=1e-2)
    gamma, lamb, clip = 0.99, 0.95, 0.2
    history = []
    for it in range(iters):
        batch = sampler.sample_batch(n_envs, auto_reset=True, record=True)      # [T, N, .] on the device
        stats = sampler.statistics()
        history.append(stats["avg_return"])
        mask = postprocess.step_mask(batch)
        obs = batch["obs"][mask].float()
        act = batch["act"][mask].float()
        values = postprocess.evaluate_values(batch, vfcn)
        adv, ret = postprocess.gae(batch, values, gamma, lamb)
        adv, ret = adv[mask], ret[mask]
        adv = (adv - adv.mean()) / (adv.std() + 1e-8)
        with torch.no_grad():
            logp_old = expl.action_dist_at(policy(obs)).log_prob(act).sum(-1)
        idx = torch.randperm(obs.shape[0], device=dev)
        for mb in idx.split(1 << 15)[:8]:
            logp = expl.action_dist_at(policy(obs[mb])).log_prob(act[mb]).sum(-1)
            ratio = torch.exp(logp - logp_old[mb])
            loss_pi = -torch.min(ratio * adv[mb], torch.clamp(ratio, 1 - clip, 1 + clip) * adv[mb]).mean()
            opt_pi.zero_grad()
            loss_pi.backward()
            opt_pi.step()
            loss_v = (vfcn(obs[mb]).squeeze(-1) - ret[mb]).pow(2).mean()
            opt_v.zero_grad()
            loss_v.backward()
            opt_v.step()
        if verbose:
            print(f"iter {it:3d}  avg return {stats['avg_return']:10.3f}  episodes {stats['num_rollouts']:7d}  "
                  f"steps {stats['num_steps']:9d}  expl std {float(expl.std.detach().mean()):.3f}")
    return history


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--envs", type=int, default=4096)
    a = ap.parse_args()
    h = train(a.iters, a.envs)
    print(f"avg return: first {h[0]:.3f} -> last {h[-1]:.3f}")
