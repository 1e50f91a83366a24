"""
GPU tests of the on-device rollout post-processing (sre_returns; SURVEY.md 8(f) rank 3;
pyrado/sampling/step_sequence.py:1049-1116): against the reference's golden values, against the oracle on random
multi-episode buffers (truncated / idle rows included), on a real record buffer through the public API, and through
size-independent properties at a large size.
"""
import numpy as np
import pytest
import torch

from conftest import golden
from simurlacra_b200 import _lib as L
from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer
from simurlacra_b200.environment_wrappers import ActNormWrapper, DomainRandWrapperLive
from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
from simurlacra_b200.policies import FNNPolicy
from simurlacra_b200.sampling import ParallelRolloutSampler, StepSequence, discounted_value, gae_returns, postprocess
from test_gpu_engine import DEV, _t

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


def test_returns_match_reference_golden(E):
    g = golden("returns_reference")
    for c in range(int(g["n_cases"])):
        r, v = g[f"c{c}_rew"], g[f"c{c}_val"]
        gamma, lamb = g[f"c{c}_gamma_lamb"]
        n = len(r)
        flag = torch.tensor([L.ROW_STEP] * (n - 1) + [L.ROW_DONE, L.ROW_TERMINAL], dtype=torch.uint8, device=DEV)[:, None]
        ret, adv = E.returns(_t(np.append(r, 0.0)[:, None], torch.float64), flag, gamma, values=_t(v[:, None], torch.float64),
                             lamb=lamb)
        np.testing.assert_allclose(ret[:n, 0].cpu().numpy(), g[f"c{c}_ret"], rtol=1e-12, atol=1e-13)
        np.testing.assert_allclose(adv[:n, 0].cpu().numpy(), g[f"c{c}_adv"], rtol=1e-12, atol=1e-13)
        ret32, adv32 = E.returns(_t(np.append(r, 0.0)[:, None], torch.float32), flag, gamma,
                                 values=_t(v[:, None], torch.float32), lamb=lamb)
        np.testing.assert_allclose(ret32[:n, 0].cpu().numpy(), g[f"c{c}_ret"], rtol=2e-4, atol=2e-4)
        np.testing.assert_allclose(adv32[:n, 0].cpu().numpy(), g[f"c{c}_adv"], rtol=2e-4, atol=2e-4)


def _random_buffer(T, N, seed):
    """Random episode layouts per lane: runs of steps closed by DONE + TERMINAL, a truncated tail or idle rows."""
    rng = np.random.RandomState(seed)
    flag = np.full((T, N), L.ROW_IDLE, dtype=np.uint8)
    for n in range(N):
        t = 0
        single = rng.rand() < 0.3
        while t < T:
            ln = int(rng.randint(1, max(2, T // 2)))
            if t + ln + 1 <= T:
                flag[t:t + ln, n] = L.ROW_STEP
                flag[t + ln - 1, n] = L.ROW_DONE
                flag[t + ln, n] = L.ROW_TERMINAL
                t += ln + 1
                if single:
                    break
            else:
                if T - t >= 2:
                    flag[t:T - 1, n] = L.ROW_STEP
                    flag[T - 1, n] = L.ROW_TRUNCATED
                elif T - t == 1:
                    flag[t, n] = L.ROW_TRUNCATED
                break
    return rng.randn(T, N), flag, rng.randn(T, N)


@pytest.mark.parametrize("T,N", [(1, 1), (2, 3), (7, 33), (61, 130), (200, 257)])
def test_returns_match_oracle_on_random_layouts(E, orc, T, N):
    rew, flag, val = _random_buffer(T, N, seed=T * 1000 + N)
    for gamma, lamb in ((0.99, 0.95), (1.0, 1.0), (0.0, 0.0)):
        r0, a0 = orc.returns(rew, flag, gamma, values=val, lamb=lamb)
        r1, a1 = E.returns(_t(rew, torch.float64), _t(flag, torch.uint8), gamma, values=_t(val, torch.float64), lamb=lamb)
        np.testing.assert_allclose(r1.cpu().numpy(), r0, rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(a1.cpu().numpy(), a0, rtol=1e-12, atol=1e-12)
        r2, none = E.returns(_t(rew, torch.float64), _t(flag, torch.uint8), gamma)
        assert none is None
        np.testing.assert_allclose(r2.cpu().numpy(), orc.returns(rew, flag, gamma)[0], rtol=1e-12, atol=1e-12)
        r3, a3 = E.returns(_t(rew, torch.float32), _t(flag, torch.uint8), gamma, values=_t(val, torch.float32), lamb=lamb)
        np.testing.assert_allclose(r3.cpu().numpy(), r0, rtol=1e-3, atol=2e-3)
        np.testing.assert_allclose(a3.cpu().numpy(), a0, rtol=1e-3, atol=2e-3)


def test_returns_argument_errors(E):
    rew = torch.zeros(4, 8, device=DEV)
    flag = torch.zeros(4, 8, dtype=torch.uint8, device=DEV)
    with pytest.raises(L.EngineError):
        E.returns(rew, flag, 1.5)
    with pytest.raises(ValueError):
        E.returns(rew, flag[:2], 0.9)
    with pytest.raises(ValueError):
        E.returns(rew.cpu(), flag.cpu(), 0.9)
    assert E.returns(torch.zeros(0, 8, device=DEV), torch.zeros(0, 8, dtype=torch.uint8, device=DEV), 0.9)[0].shape == (0, 8)


def test_postprocess_on_sampled_batch_equals_per_rollout_host_functions():
    """sample_batch(record=True) -> postprocess.gae on the strided `rew` view of the record buffer; every complete
    episode of a few lanes is cut out as a StepSequence and run through the host functions that mirror the reference."""
    env = QQubeSwingUpSim(dt=1 / 250.0, max_steps=60)
    env = DomainRandWrapperLive(ActNormWrapper(env), create_default_randomizer(env))
    torch.manual_seed(3)
    pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
    pol.param_values = pol.param_values * 8  # large actions: episodes fail early, several per lane
    vf = torch.nn.Sequential(torch.nn.Linear(6, 8), torch.nn.Tanh(), torch.nn.Linear(8, 1)).to(DEV)
    s = ParallelRolloutSampler(env, pol, 1, min_rollouts=1, seed=0)
    batch = s.sample_batch(512, auto_reset=True, record=True)
    assert not batch["rew"].is_contiguous()  # a column of the record buffer
    values = postprocess.evaluate_values(batch, vf, chunk_rows=16)
    adv, ret = postprocess.gae(batch, values, gamma=0.98, lamb=0.9)
    only = postprocess.discounted_returns(batch, 0.98)
    assert torch.equal(only, ret) or torch.allclose(only[postprocess.step_mask(batch)], ret[postprocess.step_mask(batch)])
    flag = batch["row_flag"].cpu().numpy()
    rew, val = batch["rew"].double().cpu().numpy(), values.double().cpu().numpy()
    n_checked = 0
    for n in range(0, 512, 37):
        t = 0
        while t < flag.shape[0]:
            if flag[t, n] > L.ROW_DONE:
                t += 1
                continue
            b = t
            while flag[b, n] <= L.ROW_DONE:
                b += 1
            if flag[b, n] == L.ROW_TERMINAL:
                ro = StepSequence(rewards=rew[t:b, n], observations=np.zeros((b - t + 1, 1)), actions=np.zeros((b - t, 1)),
                                  values=val[t:b + 1, n], complete=True)
                np.testing.assert_allclose(ret[t:b, n].cpu().numpy(), np.asarray(discounted_value(ro, 0.98)).reshape(-1),
                                           rtol=1e-4, atol=1e-5)
                np.testing.assert_allclose(adv[t:b, n].cpu().numpy(), np.asarray(gae_returns(ro, 0.98, 0.9)).reshape(-1),
                                           rtol=1e-4, atol=1e-4)
                n_checked += 1
            t = b + 1
    assert n_checked >= 6
    assert float(ret[~postprocess.step_mask(batch)].abs().max()) == 0.0


def test_returns_properties_at_scale(E):
    """2^17 lanes x 301 rows (fp32): linear in the rewards for fixed flags, and with gamma = 1 the first row of every
    episode holds the episode's reward sum."""
    T, N = 301, 1 << 17
    g = torch.Generator(device=DEV).manual_seed(0)
    ln = torch.randint(5, 150, (N,), device=DEV, generator=g)
    rows = torch.arange(T, device=DEV)[:, None]
    # per lane: episode A of length ln, closing row, episode B up to the end (truncated)
    flag = torch.full((T, N), L.ROW_STEP, dtype=torch.uint8, device=DEV)
    flag[rows == (ln - 1)[None, :]] = L.ROW_DONE
    flag[rows == ln[None, :]] = L.ROW_TERMINAL
    flag[T - 1] = L.ROW_TRUNCATED
    r1 = torch.randn(T, N, device=DEV, generator=g)
    r2 = torch.randn(T, N, device=DEV, generator=g)
    a = E.returns(r1, flag, 0.97)[0]
    b = E.returns(r2, flag, 0.97)[0]
    c = E.returns(2.5 * r1 + r2, flag, 0.97)[0]
    assert torch.allclose(c, 2.5 * a + b, rtol=1e-4, atol=1e-4)
    s = E.returns(r1, flag, 1.0)[0]
    first = (r1 * (rows < ln[None, :])).sum(dim=0)
    assert torch.allclose(s[0], first, rtol=1e-4, atol=1e-3)
    second = (r1 * ((rows > ln[None, :]) & (rows < T - 1))).sum(dim=0)
    assert torch.allclose(s.gather(0, (ln + 1)[None, :])[0], second, rtol=1e-4, atol=1e-3)
