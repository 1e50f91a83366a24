"""
GPU tests of the step-path wrappers fused after the first round of rows:
ObsNormWrapper (pyrado/environment_wrappers/observation_normalization.py:41-128) and GaussianActNoiseWrapper
(action_noise.py:38-102) -- kernels against the oracle on the same Philox streams, the fused rollout against the
stepping loop through the public API, and the statistics of the injected noise.
"""
import numpy as np
import pytest
import torch

from simurlacra_b200 import _lib as L
from simurlacra_b200.environment_wrappers import (ActNormWrapper, EnvWrapperObs, GaussianActNoiseWrapper,
                                                   GaussianObsNoiseWrapper, ObsNormWrapper)
from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
from simurlacra_b200.policies import FNNPolicy, IdlePolicy
from simurlacra_b200.sampling import ParallelRolloutSampler, rollout
from simurlacra_b200.sampling.rollout import compile_chain
from test_gpu_engine import DEV, _t

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


@pytest.mark.parametrize("name,dt", [("qq-su", 1 / 250.0), ("qbb", 0.01), ("qcp-su", 1 / 250.0)])
@pytest.mark.parametrize("pos", [L.ENV_ACT_NOISE_OUTER, L.ENV_ACT_NOISE_INNER])
def test_obs_norm_and_env_act_noise_match_oracle(E, orc, name, dt, pos):
    """fp64 build, injected actions (no float32 policy in the loop): observation noise -> normalisation and the action
    noise outside / inside ActNorm against orc_rollout_batch on the same Philox streams, to fp64 round-off."""
    kind = E.KIND_BY_NAME[name]
    d = E.dims(kind)
    N, H = 48, 80
    rng = np.random.RandomState(3)
    lb = -np.abs(rng.rand(d["do"]) * 3 + 1)
    ub = np.abs(rng.rand(d["do"]) * 3 + 1)
    std, mean = np.array([0.2, 0.1])[: d["da"]], np.array([0.05, -0.02])[: d["da"]]
    acts = rng.uniform(-0.5, 0.5, size=(H, N, d["da"]))
    kw = dict(act_norm=True, noise_std=0.02 * np.ones(d["do"]), noise_mean=0.01 * np.ones(d["do"]))
    cfg = E.make_cfg(kind, dt, H, dtype=torch.float64, obs_norm=(lb, ub), env_act_noise=(pos, std, mean), **kw)
    res = E.rollout(cfg, E.make_policy(L.POL_ACTIONS), N, H + 1, device=DEV, actions=_t(acts, torch.float64), seed=9,
                    env_offset=7)
    ref = orc.rollout_batch(kind, dt, H, N, H, policy=orc.make_policy(orc.POL_ACTIONS), actions=acts, seed=9, env_offset=7,
                            obs_norm=(lb, ub), env_act_noise=(pos, std, mean), **kw)
    ln = res["length"].cpu().numpy()
    np.testing.assert_array_equal(ln, ref["length"])
    st, ob, rw = (res[k].cpu().numpy() for k in ("state", "obs", "rew"))
    for n in range(N):
        T = ln[n]
        np.testing.assert_allclose(st[: T + 1, n], ref["state"][: T + 1, n], rtol=1e-9, atol=1e-10)
        np.testing.assert_allclose(ob[: T + 1, n], ref["obs"][: T + 1, n], rtol=1e-9, atol=1e-10)
        np.testing.assert_allclose(rw[:T, n], ref["rew"][:T, n], rtol=1e-9, atol=1e-10)
    # the wrappers really act: other position / no normalisation give other trajectories
    plain = E.rollout(E.make_cfg(kind, dt, H, dtype=torch.float64, **kw), E.make_policy(L.POL_ACTIONS), N, H + 1, device=DEV,
                      actions=_t(acts, torch.float64), seed=9, env_offset=7)
    assert not np.allclose(plain["state"][1].cpu().numpy(), st[1])
    assert not np.allclose(plain["obs"][0].cpu().numpy(), ob[0])
    # float build tracks the fp64 build
    cfg32 = E.make_cfg(kind, dt, H, dtype=torch.float32, obs_norm=(lb, ub), env_act_noise=(pos, std, mean), **kw)
    r32 = E.rollout(cfg32, E.make_policy(L.POL_ACTIONS), N, H + 1, device=DEV, actions=_t(acts, torch.float32), seed=9,
                    env_offset=7)
    np.testing.assert_allclose(r32["obs"][:5].double().cpu().numpy(), ob[:5], rtol=2e-4, atol=2e-4)


def test_cfg_argument_errors(E):
    with pytest.raises(L.EngineError):
        E.rollout(E.make_cfg(L.QQ, 0.01, 5, obs_norm=(np.zeros(6), np.zeros(6))), E.make_policy(L.POL_IDLE), 2, 6, device=DEV)
    with pytest.raises(L.EngineError):
        E.rollout(E.make_cfg(L.QQ, 0.01, 5, obs_norm=(-np.ones(6), np.full(6, np.inf))), E.make_policy(L.POL_IDLE), 2, 6,
                  device=DEV)
    with pytest.raises(L.EngineError):
        E.rollout(E.make_cfg(L.QQ, 0.01, 5, env_act_noise=(5, [0.1], [0.0])), E.make_policy(L.POL_IDLE), 2, 6, device=DEV)


class _Transparent(EnvWrapperObs):
    """An observation wrapper the kernels do not know: forces rollout() onto the generic stepping path."""

    def _process_obs(self, obs):
        return obs


def _chain(order, max_steps=60):
    sim = QQubeSwingUpSim(dt=1 / 250.0, max_steps=max_steps)
    if order == "noise_outside_norm":
        env = GaussianActNoiseWrapper(ActNormWrapper(sim), noise_std=[0.3], noise_mean=[0.1])
    else:
        env = ActNormWrapper(GaussianActNoiseWrapper(sim, noise_std=[0.5]))
    env = GaussianObsNoiseWrapper(env, noise_std=0.01 * np.ones(6))
    return ObsNormWrapper(env)


@pytest.mark.parametrize("order", ["noise_outside_norm", "noise_inside_norm"])
def test_fused_rollout_equals_stepping_loop_with_wrappers(order):
    """One kernel launch == the loop over env.step for a chain with all four fused wrappers (same Philox keys in both
    paths); IdlePolicy, so fp64 round-off only.  The observation space is the normalised one."""
    ros = []
    for generic in (False, True):
        env = _chain(order)
        assert compile_chain(env) is not None
        np.testing.assert_array_equal(env.obs_space.bound_up, np.ones(6))
        e2 = _Transparent(env) if generic else env
        ros.append(rollout(e2, IdlePolicy(env.spec), eval=True, seed=2, sub_seed=1, sub_sub_seed=5))
    a, b = ros
    assert len(a) == len(b) and len(a) > 5
    np.testing.assert_allclose(a.states, b.states, rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(a.observations, b.observations, rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(a.rewards, b.rewards, rtol=1e-12, atol=1e-13)
    # normalised velocities: obs[4] = theta_dot / (20 pi) up to the observation noise
    np.testing.assert_allclose(a.observations[:, 4] * 20 * np.pi, a.states[:, 2], atol=0.01 * 6 + 1e-9)
    assert np.abs(a.states[1:, 2]).max() > 0  # the action noise moves the idle system
    dp = a.rollout_info["domain_param"]
    assert "act_noise_std" in dp and "obs_noise_std" in dp


def test_action_noise_statistics_through_sampler():
    """Zero policy on the ball balancer: the env sees exactly the action noise.  With ActNorm outside, the first
    step's state change is linear in the applied voltage, so its spread over many rollouts follows the noise's std."""
    std = np.array([0.2, 0.4])
    sim = QBallBalancerSim(dt=0.01, max_steps=3, num_envs=1)
    env = GaussianActNoiseWrapper(ActNormWrapper(sim), noise_std=std)
    pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
    pol.param_values = pol.param_values * 0
    s = ParallelRolloutSampler(env, pol, 1, min_rollouts=1, seed=0)
    batch = s.sample_batch(1 << 14, auto_reset=False, record=True)
    st = batch["state"].double()
    dv = (st[1, :, 4:6] - st[0, :, 4:6])  # servo velocity change of the first step ~ gain * voltage
    ratio = float(dv[:, 1].std() / dv[:, 0].std())
    assert abs(ratio - std[1] / std[0]) < 0.1
    assert abs(float(dv[:, 0].mean())) < 0.05 * float(dv[:, 0].std()) + 1e-9
    again = s.sample_batch(1 << 14, auto_reset=False, record=True)
    assert not torch.equal(again["state"][1], batch["state"][1])  # next sample() call: other keys
