"""
GPU tests of step-based exploration fused into the rollout kernel (SURVEY.md 8(f) rank 1;
pyrado/exploration/stochastic_action.py:68-94, normal_noise.py:155-163, uniform_noise.py:160-168):
the kernel's explorative actions against the oracle on the same Philox stream, their statistics, determinism /
shard invariance, the tensor-core policy build, and the reference-facing API (NormalActNoiseExplStrat around a policy,
through rollout() and ParallelRolloutSampler, evaluate() -> log_prob for the learner).
"""
import numpy as np
import pytest
import torch

from simurlacra_b200 import _lib as L
from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer
from simurlacra_b200.environment_wrappers import ActNormWrapper, DomainRandWrapperLive
from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
from simurlacra_b200.exploration import NormalActNoiseExplStrat, UniformActNoiseExplStrat
from simurlacra_b200.policies import FNNPolicy
from simurlacra_b200.sampling import ParallelRolloutSampler, rollout
from test_gpu_engine import DEV, _t

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


def _weights(sizes, seed=3, scale=0.4):
    rng = np.random.RandomState(seed)
    return (rng.randn(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1))) * scale).astype(np.float32)


@pytest.mark.parametrize("name,dt,da", [("qq-su", 1 / 250.0, 1), ("qbb", 0.01, 2)])
@pytest.mark.parametrize("noise", ["normal", "uniform"])
def test_explorative_rollout_matches_oracle(E, orc, name, dt, da, noise):
    """fp64 build, FNN[16,16] + exploration noise in the loop against orc_rollout_batch on the same Philox stream: the
    noise is float32 on both sides (fp64 Box-Muller rounded to float in the oracle, float32 math on the device), so the
    closed loop agrees to the float32-policy tolerance (1e-4) and the episode lengths are identical."""
    kind = E.KIND_BY_NAME[name]
    d = E.dims(kind)
    N, H = 64, 100
    sizes = [d["do"], 16, 16, da]
    w = _weights(sizes)
    k = L.ACT_NOISE_NORMAL if noise == "normal" else L.ACT_NOISE_UNIFORM
    scale, mean = [0.3, 0.2][:da], [0.05, -0.1][:da]
    cfg = E.make_cfg(kind, dt, H, dtype=torch.float64, act_norm=True)
    pol = E.make_policy(L.POL_FNN, _t(w, torch.float32), sizes=sizes, act_noise=k, act_noise_scale=scale, act_noise_mean=mean)
    res = E.rollout(cfg, pol, N, H + 1, device=DEV, seed=21, env_offset=5)
    opol = orc.make_policy(orc.POL_FNN, w, sizes=sizes, act_noise=k, act_noise_scale=scale, act_noise_mean=mean)
    ref = orc.rollout_batch(kind, dt, H, N, H, policy=opol, act_norm=True, seed=21, env_offset=5)
    ln = res["length"].cpu().numpy()
    np.testing.assert_array_equal(ln, ref["length"])
    st, ac = res["state"].cpu().numpy(), res["act"].cpu().numpy()
    for n in range(N):
        T = ln[n]
        sc = np.maximum(1.0, np.abs(ref["state"][: T + 1, n]))
        assert (np.abs(st[: T + 1, n] - ref["state"][: T + 1, n]) / sc).max() < 1e-4
        assert np.abs(ac[:T, n] - ref["act"][:T, n]).max() < 1e-4
    # first step: same observation on both sides, so the explorative action agrees to float32 round-off of the noise
    np.testing.assert_allclose(ac[0], ref["act"][0], rtol=0, atol=2e-6)
    # and the noise is really there
    quiet = E.rollout(cfg, E.make_policy(L.POL_FNN, _t(w, torch.float32), sizes=sizes), N, H + 1, device=DEV, seed=21,
                      env_offset=5)
    assert np.abs(quiet["act"].cpu().numpy()[0] - ac[0]).min() > 1e-6


@pytest.mark.parametrize("prec", [L.PREC_FP32, L.PREC_TF32X3])
def test_noise_statistics_and_determinism(E, prec):
    """Zero policy (all weights 0): the recorded action IS mean + std * eps.  Moments over 2^15 envs x 60 steps, all
    draws distinct across envs and steps, same seed -> same bits, shard offset -> same lanes; fp32 FMA and tcgen05 builds."""
    N, H = 1 << 15, 60
    sizes = [6, 64, 64, 1]
    w = torch.zeros(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(3)), dtype=torch.float32, device=DEV)
    cfg = E.make_cfg(L.QQ, 1 / 250.0, H, dtype=torch.float32, act_norm=True)
    pol = E.make_policy(L.POL_FNN, w, sizes=sizes, precision=prec, act_noise=L.ACT_NOISE_NORMAL, act_noise_scale=[0.25],
                        act_noise_mean=[0.1])
    a = E.rollout(cfg, pol, N, H + 1, device=DEV, seed=4, stop_on_done=False)
    act = a["act"][:H, :, 0].double()
    assert abs(float(act.mean()) - 0.1) < 1e-3 and abs(float(act.std()) - 0.25) < 1e-3
    z = (act - 0.1) / 0.25
    assert abs(float((z ** 3).mean())) < 0.02 and abs(float((z ** 4).mean()) - 3.0) < 0.06  # skewness, kurtosis
    assert float(z.abs().max()) < 5.8  # the engine's normal is bounded by sqrt(48 ln 2)
    assert len(torch.unique(act)) > 0.95 * act.numel()
    c = torch.corrcoef(torch.stack([act[:-1].reshape(-1), act[1:].reshape(-1)]))[0, 1]
    assert abs(float(c)) < 0.01  # consecutive steps are independent draws
    b = E.rollout(cfg, pol, N, H + 1, device=DEV, seed=4, stop_on_done=False)
    assert torch.equal(a["act"], b["act"]) and torch.equal(a["state"], b["state"])
    sub = E.rollout(cfg, pol, 256, H + 1, device=DEV, seed=4, env_offset=1000, stop_on_done=False)
    assert torch.equal(sub["act"], a["act"][:, 1000:1256])
    u = E.make_policy(L.POL_FNN, w, sizes=sizes, precision=prec, act_noise=L.ACT_NOISE_UNIFORM, act_noise_scale=[0.5])
    ua = E.rollout(cfg, u, N, H + 1, device=DEV, seed=4, stop_on_done=False)["act"][:H, :, 0].double()
    assert float(ua.min()) >= -0.5 and float(ua.max()) < 0.5 and abs(float(ua.std()) - 0.5 / np.sqrt(3)) < 2e-3


def test_noise_argument_errors(E):
    w = torch.zeros(6 * 8 + 8 + 8 + 1, dtype=torch.float32, device=DEV)
    cfg = E.make_cfg(L.QQ, 1 / 250.0, 10, dtype=torch.float32)
    with pytest.raises(L.EngineError):
        E.rollout(cfg, E.make_policy(L.POL_FNN, w, sizes=[6, 8, 1], act_noise=L.ACT_NOISE_NORMAL, act_noise_scale=[0.0]),
                  4, 11, device=DEV)
    with pytest.raises(L.EngineError):
        E.rollout(cfg, E.make_policy(L.POL_FNN, w, sizes=[6, 8, 1], act_noise=7, act_noise_scale=[0.1]), 4, 11, device=DEV)
    with pytest.raises(L.EngineError):
        E.rollout(cfg, E.make_policy(L.POL_IDLE, act_noise=L.ACT_NOISE_NORMAL, act_noise_scale=[0.1]), 4, 11, device=DEV)


def _chain(max_steps=80):
    env = QQubeSwingUpSim(dt=1 / 250.0, max_steps=max_steps)
    return DomainRandWrapperLive(ActNormWrapper(env), create_default_randomizer(env))


def test_expl_strat_through_rollout_and_sampler():
    """The reference's usage (Pyrado/tests/test_sampling.py:570-646 with an exploration strategy as the policy):
    sampler == sequential rollout() for every index, explorative actions scatter around the wrapped policy's output
    with the strategy's std, and evaluate() gives the learner log-probabilities with gradients."""
    env = _chain()
    torch.manual_seed(1)
    pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
    ex = NormalActNoiseExplStrat(pol, std_init=0.2)
    n = 12
    ros = ParallelRolloutSampler(env, ex, num_workers=2, min_rollouts=n, seed=3).sample()
    seq = [rollout(env, ex, seed=3, sub_seed=0, sub_sub_seed=i) for i in range(n)]
    det = ParallelRolloutSampler(env, pol, num_workers=2, min_rollouts=n, seed=3).sample()
    for a, b, c in zip(ros, seq, det):
        assert len(a) == len(b)
        np.testing.assert_array_equal(a.actions, b.actions)
        np.testing.assert_array_equal(a.states, b.states)
        assert a.rollout_info["domain_param"] == c.rollout_info["domain_param"]  # same DR draw, other stream
        assert not np.allclose(a.actions[: min(len(a), len(c))], c.actions[: min(len(a), len(c))])
    # residuals around the wrapped policy: (act - policy(obs)) / std ~ N(0, 1)
    obs = torch.as_tensor(np.concatenate([r.observations[:-1] for r in ros]), dtype=torch.float32)
    act = torch.as_tensor(np.concatenate([r.actions for r in ros]), dtype=torch.float32)
    with torch.no_grad():
        z = (act - pol(obs)) / 0.2
    assert abs(float(z.mean())) < 0.15 and abs(float(z.std()) - 1.0) < 0.15
    from simurlacra_b200.sampling import StepSequence

    cat = StepSequence.concat(ros, data_format="torch")
    lp = ex.evaluate(cat).log_prob(cat.actions.to(torch.float32))
    assert lp.shape == cat.actions.shape and torch.isfinite(lp).all()
    lp.sum().backward()
    assert ex.noise.log_std.grad is not None and float(pol.param_grad.abs().sum()) > 0
    # uniform strategy, two action dimensions, tensor-core policy build
    env2 = ActNormWrapper(QBallBalancerSim(dt=0.01, max_steps=50))
    torch.manual_seed(2)
    pol2 = FNNPolicy(env2.spec, hidden_sizes=[64, 64], hidden_nonlin=torch.tanh)
    un = UniformActNoiseExplStrat(pol2, halfspan_init=0.1)
    un.engine_precision = L.PREC_TF32X3
    ros2 = ParallelRolloutSampler(env2, un, 1, min_rollouts=6, seed=0).sample()
    for r in ros2:
        with torch.no_grad():
            res = torch.as_tensor(r.actions, dtype=torch.float32) - pol2(torch.as_tensor(r.observations[:-1], dtype=torch.float32))
        assert float(res.abs().max()) < 0.1 + 1e-4 and float(res.abs().max()) > 0.05
