"""
GPU tests of the stateful / shape-changing step-path wrappers: ActDelayWrapper,
DownsamplingWrapper, ObsPartialWrapper, ObsVelFiltWrapper, StateAugmentationWrapper
(pyrado/environment_wrappers/action_delay.py, downsampling.py, observation_partial.py, observation_velfilter.py,
state_augmentation.py) around QQubeSwingUpSim stepping on the step kernels, against trajectories of the UNMODIFIED
reference wrappers (tests/golden/wrappers_reference.npz, made by `oracle/gen_golden.py wrappers`): single env in fp64
like the reference, a float32 batch with per-env data, through rollout() / ParallelRolloutSampler, and -- the fused path
-- inside the persistent rollout kernel (one wrapper state per lane, sre_step_wrappers) against the same trajectories.
"""
import numpy as np
import pytest
import torch

from conftest import golden
from simurlacra_b200.environment_wrappers import (ActDelayWrapper, ActNormWrapper, DownsamplingWrapper, ObsPartialWrapper,
                                                   ObsVelFiltWrapper, StateAugmentationWrapper)
from simurlacra_b200.environment_wrappers.utils import inner_env
from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
from simurlacra_b200.policies import FNNPolicy
from simurlacra_b200.sampling import ParallelRolloutSampler, rollout
from simurlacra_b200.sampling.rollout import compile_chain

pytestmark = pytest.mark.gpu

CASES = dict(
    delay3=lambda e: ActDelayWrapper(e, delay=3),
    delay3_norm=lambda e: ActNormWrapper(ActDelayWrapper(e, delay=3)),
    down4=lambda e: DownsamplingWrapper(e, factor=4),
    partial_mask=lambda e: ObsPartialWrapper(e, mask=[0, 0, 0, 0, 1, 1]),
    partial_keep=lambda e: ObsPartialWrapper(e, idcs=["sin_theta", "cos_theta", "theta_dot"], keep_selected=True),
    velfilt=lambda e: ObsVelFiltWrapper(e, idcs_pos=["theta", "alpha"], idcs_vel=["theta_dot", "alpha_dot"]),
    augment=lambda e: StateAugmentationWrapper(e, domain_param=["mass_pend_pole", "length_pend_pole"]),
    augment_fixed=lambda e: StateAugmentationWrapper(e, domain_param=["mass_pend_pole", "length_pend_pole"], fixed=True),
)


def _inputs(g, name):
    dp = {str(k): float(v) for k, v in zip(g["dp_keys"], g["dp_vals"])}
    act = g["act"] / 5.0 if name.endswith("_norm") else g["act"]
    return dp, act, int(g[name + "_len"])


@pytest.mark.parametrize("name", list(CASES))
def test_single_env_matches_reference(name):
    g = golden("wrappers_reference")
    dp, act, T = _inputs(g, name)
    env = CASES[name](QQubeSwingUpSim(dt=1 / 250.0, max_steps=60))
    np.testing.assert_allclose(env.obs_space.bound_lo, g[name + "_obs_lo"], rtol=1e-12)
    np.testing.assert_allclose(env.obs_space.bound_up, g[name + "_obs_up"], rtol=1e-12)
    assert compile_chain(env).fusion.wrap  # env.step runs them between step launches; rollouts fuse them (below)
    for episode in range(2):  # the second reset starts from clean wrapper state (queue, counter, filter)
        obs = [env.reset(init_state=g["init_state"].copy(), domain_param=dict(dp))]
        states, rews = [env.state], []
        for t in range(T):
            o, r, d, _ = env.step(act[t].copy())
            obs.append(o)
            states.append(env.state)
            rews.append(float(r))
        if name == "down4" and episode == 1:
            # reference behaviour (downsampling.py:118-131): the first observation after a reset is filtered together
            # with what the previous episode left in the buffer
            assert not np.allclose(obs[0], g[name + "_obs"][0])
            obs, ref_obs = obs[4:], g[name + "_obs"][4:T + 1]
        else:
            ref_obs = g[name + "_obs"][: T + 1]
        np.testing.assert_allclose(np.array(obs), ref_obs, rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(np.array(states), g[name + "_states"][: T + 1], rtol=1e-9, atol=1e-10)
        np.testing.assert_allclose(np.array(rews), g[name + "_rews"][:T], rtol=1e-9, atol=1e-10)


@pytest.mark.parametrize("name", list(CASES))
def test_batch_env_matches_reference(name):
    """8 envs in float32 on the device, every env fed the golden inputs: all rows follow the reference to float32
    round-off; tensors in, tensors out."""
    g = golden("wrappers_reference")
    dp, act, T = _inputs(g, name)
    n = 8
    env = CASES[name](QQubeSwingUpSim(dt=1 / 250.0, max_steps=60, num_envs=n))
    dev = inner_env(env).device
    o = env.reset(init_state=np.tile(g["init_state"], (n, 1)), domain_param=dict(dp))
    assert isinstance(o, torch.Tensor) and o.shape == (n, g[name + "_obs"].shape[1])
    obs = [o]
    for t in range(T):
        o, r, d, _ = env.step(torch.as_tensor(np.tile(act[t], (n, 1)), dtype=torch.float32, device=dev))
        obs.append(o)
    obs = torch.stack(obs).double().cpu().numpy()
    for i in range(n):
        np.testing.assert_allclose(obs[:, i], g[name + "_obs"][: T + 1], rtol=2e-3, atol=2e-3)


def test_per_env_data_in_a_batch():
    """What only a batch has: one delay queue / filter state / parameter row per env."""
    n = 4
    sim = QQubeSwingUpSim(dt=1 / 250.0, max_steps=30, num_envs=n)
    env = StateAugmentationWrapper(sim, domain_param=["mass_pend_pole"])
    masses = np.array([0.02, 0.024, 0.028, 0.032])
    o = env.reset(domain_param=dict(mass_pend_pole=masses))
    np.testing.assert_allclose(o[:, -1].double().cpu().numpy(), masses, rtol=1e-6)
    assert env.offset == 6 and env.mask.tolist() == [0] * 6 + [1]
    # delay: env i gets its own action sequence, each delayed by 2 steps
    env = ActDelayWrapper(QQubeSwingUpSim(dt=1 / 250.0, max_steps=30, num_envs=n), delay=2)
    ref = QQubeSwingUpSim(dt=1 / 250.0, max_steps=30, num_envs=n)
    s0 = np.tile(np.array([0.1, 0.2, 0.0, 0.0]), (n, 1))
    env.reset(init_state=s0)
    ref.reset(init_state=s0)
    acts = torch.linspace(-2, 2, 12 * n, device=ref.device).reshape(12, n, 1)
    for t in range(12):
        env.step(acts[t])
        ref.step(acts[t - 2] if t >= 2 else torch.zeros(n, 1, device=ref.device))
    assert torch.equal(env.state, ref.state)
    assert env.domain_param["act_delay"] == 2
    env.domain_param = dict(act_delay=5)
    assert env.delay == 5
    # velocity filter: steady state at reset -> zero velocity estimates, whatever the initial position
    vf = ObsVelFiltWrapper(QQubeSwingUpSim(dt=1 / 250.0, max_steps=30, num_envs=n), idcs_pos=["theta", "alpha"],
                           idcs_vel=["theta_dot", "alpha_dot"])
    o = vf.reset(init_state=np.tile(np.array([0.3, -0.2, 1.0, -1.0]), (n, 1)))
    assert float(o[:, 4:].abs().max()) < 1e-5 and vf.z.shape == (1, n, 2)


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-9), (torch.float32, 2e-3)])
def test_fused_in_the_persistent_kernel_matches_reference(name, dtype, tol):
    """The whole horizon in ONE sre_rollout launch with the wrapper state per lane (injected golden actions): the
    recorded observations are the wrapper's, states and rewards follow the reference."""
    import simurlacra_b200.engine as E
    from simurlacra_b200 import _lib as L

    g = golden("wrappers_reference")
    dp, act, T = _inputs(g, name)
    n = 5
    env = CASES[name](QQubeSwingUpSim(dt=1 / 250.0, max_steps=60, num_envs=n, dtype=dtype))
    env.domain_param = dict(dp)
    cc = compile_chain(env)
    assert cc is not None and cc.fusion.wrap
    sim = cc.sim
    cfg = sim._make_cfg(cc.fusion)
    dev = sim.device
    init = torch.as_tensor(np.tile(g["init_state"], (n, 1)), dtype=dtype, device=dev)
    acts = torch.as_tensor(np.tile(act[:T, None, :], (1, n, 1)), dtype=dtype, device=dev).contiguous()
    res = E.rollout(cfg, E.make_policy(L.POL_ACTIONS), n, T + 1, device=dev, init=init, actions=acts, seed=3)
    assert res["length"].tolist() == [T] * n
    n_rec, aug_idx, aug_nom, aug_fixed = cc.obs_layout
    obs = res["obs"][: T + 1, :, :n_rec].double().cpu().numpy()
    ref = g[name + "_obs"][: T + 1]
    for i in range(n):
        np.testing.assert_allclose(obs[:, i], ref[:, :n_rec], rtol=tol, atol=tol)
        np.testing.assert_allclose(res["state"][: T + 1, i].double().cpu().numpy(), g[name + "_states"][: T + 1], rtol=tol, atol=tol)
        np.testing.assert_allclose(res["rew"][:T, i].double().cpu().numpy(), g[name + "_rews"][:T], rtol=tol, atol=tol)
    if aug_idx:  # the appended parameters are rebuilt from the lane's parameter row
        prm = res["params"][:, aug_idx].double().cpu().numpy() if not aug_fixed else np.tile(aug_nom, (n, 1))
        np.testing.assert_allclose(prm, np.tile(ref[0, n_rec:], (n, 1)), rtol=1e-6)


@pytest.mark.parametrize("name", list(CASES) + ["down_delay_partial"])
def test_fused_rollout_equals_the_stepping_loop(name):
    """rollout() with a fused FNN policy: the persistent kernel (wrappers fused) against the reference's loop over
    env.step with the wrappers applied between step launches (record_dts=True forces it), single env in fp64."""
    make = CASES.get(name) or (lambda e: ObsPartialWrapper(DownsamplingWrapper(ActNormWrapper(ActDelayWrapper(e, delay=2)),
                                                                               factor=3), mask=[0, 1, 0, 0, 0, 0]))
    env = make(QQubeSwingUpSim(dt=1 / 250.0, max_steps=40))
    torch.manual_seed(1)
    pol = FNNPolicy(env.spec, hidden_sizes=[12, 12], hidden_nonlin=torch.tanh)
    assert compile_chain(env) is not None
    fused = rollout(env, pol, seed=5, reset_kwargs=dict(init_state=np.array([0.1, -0.2, 0.3, 0.1])))
    env2 = make(QQubeSwingUpSim(dt=1 / 250.0, max_steps=40))
    loop = rollout(env2, pol, seed=5, record_dts=True, reset_kwargs=dict(init_state=np.array([0.1, -0.2, 0.3, 0.1])))
    assert len(fused) == len(loop) and fused.observations.shape == loop.observations.shape
    np.testing.assert_allclose(fused.observations, loop.observations, rtol=1e-5, atol=1e-5)  # float32 policy on both sides
    np.testing.assert_allclose(fused.states, loop.states, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(fused.actions, loop.actions, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(fused.rewards, loop.rewards, rtol=1e-5, atol=1e-5)


def test_through_rollout_and_sampler():
    """rollout() and the sampler hand out StepSequences with the wrapper's observation width (fused into the persistent
    kernel); the recorded action stays the policy's, like the reference."""
    env = ObsPartialWrapper(ActNormWrapper(QQubeSwingUpSim(dt=1 / 250.0, max_steps=25)), idcs=["theta_dot", "alpha_dot"])
    assert compile_chain(env) is not None and env.obs_space.flat_dim == 4
    torch.manual_seed(0)
    pol = FNNPolicy(env.spec, hidden_sizes=[8], hidden_nonlin=torch.tanh)
    ro = rollout(env, pol, seed=0)
    assert ro.observations.shape == (len(ro) + 1, 4) and ro.actions.shape == (len(ro), 1)
    full = ro.states
    np.testing.assert_allclose(ro.observations[:, 0], np.sin(full[:, 0]), atol=1e-9)
    ros = ParallelRolloutSampler(env, pol, num_workers=1, min_rollouts=3, seed=1).sample()
    assert len(ros) == 3 and all(r.observations.shape[1] == 4 for r in ros)
    aug = StateAugmentationWrapper(QQubeSwingUpSim(dt=1 / 250.0, max_steps=10), domain_param=["gravity_const"])
    pol2 = FNNPolicy(aug.spec, hidden_sizes=[8], hidden_nonlin=torch.tanh)
    r2 = rollout(aug, pol2, seed=0)
    assert r2.observations.shape[1] == 7 and np.allclose(r2.observations[:, -1], 9.81)
