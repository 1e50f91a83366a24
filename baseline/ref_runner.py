"""
BENCH INFRASTRUCTURE -- times the UNMODIFIED reference (baseline/_ref/pyrado, built by baseline/make_ref.sh) on the host
cores for config C2 of BASELINE.json: QQubeSwingUpSim(dt=1/250, max_steps=600) + ActNormWrapper +
DomainRandWrapperLive(create_default_randomizer_qq()), FNN [64, 64] tanh policy on the CPU.

  B1  ParallelRolloutSampler(env, policy, num_workers = n_cores, min_rollouts = 8 * n_cores)   (CPU baseline 1)
  B2  the same with num_workers = 1
  B3  hand loop reset + 600 x step with a fixed action on one core (Pyrado/tests/test_environments.py:87-91)

Run as a subprocess of bench.py with PYTHONPATH=baseline/_ref and SRE_REF_SHIM=1 (the sampler's workers are spawned and
pick the import shim up through baseline/_ref/sitecustomize.py).  Prints one JSON object.

The fork's own FNNPolicy hard-codes the cart-pole feature transform in forward() (policies/feed_back/fnn.py:214-218) and
cannot take QQube observations; `PlainFNNPolicy` below is the upstream behaviour -- the reference's own `FNN` module
under the reference's `Policy` base class, `forward(obs) = net(obs)`.
"""
import argparse
import json
import os
import sys
import time

import torch as to

import pyrado
from pyrado.domain_randomization.default_randomizers import create_default_randomizer_qq
from pyrado.environment_wrappers.action_normalization import ActNormWrapper
from pyrado.environment_wrappers.domain_randomization import DomainRandWrapperLive
from pyrado.environments.pysim.quanser_qube import QQubeSwingUpSim
from pyrado.policies.base import Policy
from pyrado.policies.feed_back.fnn import FNN


class PlainFNNPolicy(Policy):
    name: str = "fnn_plain"

    def __init__(self, spec, hidden_sizes, hidden_nonlin, use_cuda: bool = False):
        super().__init__(spec, use_cuda)
        self.net = FNN(input_size=spec.obs_space.flat_dim, output_size=spec.act_space.flat_dim,
                       hidden_sizes=hidden_sizes, hidden_nonlin=hidden_nonlin)
        self.init_param(None)

    def init_param(self, init_values=None, **kwargs):
        if init_values is None:
            self.net.init_param(init_values, **kwargs)
        else:
            self.param_values = init_values

    def forward(self, obs):
        return self.net(obs)


def make_env(horizon):
    return DomainRandWrapperLive(ActNormWrapper(QQubeSwingUpSim(dt=1 / 250.0, max_steps=horizon)),
                                 create_default_randomizer_qq())


def run_sampler(workers, min_rollouts, repeat, horizon):
    from pyrado.sampling.parallel_rollout_sampler import ParallelRolloutSampler

    pyrado.set_seed(0)
    env = make_env(horizon)
    policy = PlainFNNPolicy(env.spec, [64, 64], to.tanh)
    t0 = time.perf_counter()
    sampler = ParallelRolloutSampler(env, policy, num_workers=workers, min_rollouts=min_rollouts, seed=0,
                                     show_progress_bar=False)
    t_init = time.perf_counter() - t0
    sampler.sample()  # warm-up
    steps, t = 0, 0.0
    for _ in range(repeat):
        t0 = time.perf_counter()
        ros = sampler.sample()
        t += time.perf_counter() - t0
        steps += sum(len(ro) for ro in ros)
    return dict(steps=steps, seconds=t, steps_per_s=steps / t, workers=workers, min_rollouts=min_rollouts, repeat=repeat,
                pool_start_s=t_init)


def run_hand_loop(episodes, horizon):
    import numpy as np

    pyrado.set_seed(0)
    env = QQubeSwingUpSim(dt=1 / 250.0, max_steps=horizon)
    act = np.array([0.1])
    steps = 0
    t0 = time.perf_counter()
    for _ in range(episodes):
        env.reset()
        done = False
        while not done:
            _, _, done, _ = env.step(act)
            steps += 1
    t = time.perf_counter() - t0
    return dict(steps=steps, seconds=t, steps_per_s=steps / t, workers=1, episodes=episodes)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="sampler", choices=["sampler", "hand"])
    ap.add_argument("--workers", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--min-rollouts", type=int, default=0)
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--episodes", type=int, default=5)
    ap.add_argument("--horizon", type=int, default=600)
    a = ap.parse_args()
    to.set_num_threads(1)
    if a.mode == "hand":
        out = run_hand_loop(a.episodes, a.horizon)
    else:
        out = run_sampler(a.workers, a.min_rollouts or 8 * a.workers, a.repeat, a.horizon)
    print("REF_RESULT " + json.dumps(out), flush=True)
