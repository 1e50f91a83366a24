"""
GPU tests of the reference-facing Python API (SimPyEnv.reset/step, wrappers, DomainRandomizer, rollout(),
ParallelRolloutSampler, step_diff_state) -- every call ends in the CUDA kernels through the C-ABI.  Numbers are checked
against the golden vectors the unmodified reference produced (tests/golden) and against the reference's behavioural
contracts (Pyrado/tests/test_environments.py, test_domain_randomization.py, test_sampling.py:570-701).
"""
import numpy as np
import pytest
import torch

import simurlacra_b200 as pyrado
from conftest import ENV_FILES, OMO_TASK, golden
from simurlacra_b200.domain_randomization import DomainRandomizer, NormalDomainParam, UniformDomainParam
from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer
from simurlacra_b200.environment_wrappers import (ActNormWrapper, DomainRandWrapperBuffer, DomainRandWrapperLive,
                                                   EnvWrapperObs, GaussianObsNoiseWrapper, ObsActCatWrapper)
from simurlacra_b200.environments.pysim.ball_on_beam import BallOnBeamDiscSim, BallOnBeamSim
from simurlacra_b200.environments.pysim.one_mass_oscillator import OneMassOscillatorSim
from simurlacra_b200.environments.pysim.pendulum import PendulumSim
from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
from simurlacra_b200.environments.pysim.quanser_cartpole import QCartPoleStabSim, QCartPoleSwingUpSim
from simurlacra_b200.environments.pysim.quanser_qube import QQubeStabSim, QQubeSwingUpSim
from simurlacra_b200.policies import FeatureStack, FNNPolicy, IdlePolicy, LinearPolicy
from simurlacra_b200.policies.features import const_feat, identity_feat, squared_feat
from simurlacra_b200.sampling import ParallelRolloutSampler, StepSequence, rollout

pytestmark = pytest.mark.gpu


def _make_env(kind_name, dt, max_steps, kw, **extra):
    if kind_name == "pend":
        return PendulumSim(dt=dt, max_steps=max_steps, **extra)
    if kw.get("stab") == "qq-st":
        return QQubeStabSim(dt=dt, max_steps=max_steps, **extra)
    if kw.get("stab") == "qcp-st":
        return QCartPoleStabSim(dt=dt, max_steps=max_steps, long=kw.get("long_pole", False),
                                simple_dynamics=kw.get("simple_dynamics", False), **extra)
    if kind_name == "qq-su":
        return QQubeSwingUpSim(dt=dt, max_steps=max_steps, **extra)
    if kind_name == "qcp-su":
        return QCartPoleSwingUpSim(dt=dt, max_steps=max_steps, long=kw.get("long_pole", False),
                                   simple_dynamics=kw.get("simple_dynamics", False), **extra)
    if kind_name == "qbb":
        return QBallBalancerSim(dt=dt, max_steps=max_steps, simple_dynamics=kw.get("simple_dynamics", False), **extra)
    if kind_name == "bob" and kw.get("disc_act"):
        return BallOnBeamDiscSim(dt=dt, max_steps=max_steps, **extra)
    if kind_name == "bob":
        return BallOnBeamSim(dt=dt, max_steps=max_steps, **extra)
    return OneMassOscillatorSim(dt=dt, max_steps=max_steps, task_args=OMO_TASK, **extra)


@pytest.mark.parametrize("fname,kind_name,kw", ENV_FILES)
def test_env_reset_step_match_reference_golden(fname, kind_name, kw):
    """SimPyEnv.reset/step with the reference's types (ndarray, float, bool, dict) reproduce the reference's
    trajectories: fp64, 1e-10 relative per step over the first 100 steps of every golden case."""
    g = golden(fname)
    for c in range(int(g["n_cases"])):
        env = _make_env(kind_name, float(g["dt"]), int(g["max_steps"]), kw)
        S, A, R, D, O = (g[f"c{c}_{k}"] for k in ("states", "actions", "rews", "dones", "obs"))
        dp = dict(zip(env._param_names, (float(v) for v in g[f"c{c}_params"])))
        obs = env.reset(init_state=S[0], domain_param=dp)
        if env._dims["nh"]:
            env._hidden[0, : env._dims["nh"]] = torch.as_tensor(g[f"c{c}_hidden"][0]).to(env._hidden)
        assert isinstance(obs, np.ndarray) and obs.dtype == np.float64 and obs.shape == env.obs_space.shape
        np.testing.assert_allclose(obs, O[0], rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(env.state, S[0], rtol=0, atol=0)
        for t in range(min(len(A), 100)):
            obs, rew, done, info = env.step(A[t].copy())
            assert isinstance(rew, float) and isinstance(done, bool) and info == dict()
            scale = np.maximum(1.0, np.abs(S[t + 1]))
            assert (np.abs(env.state - S[t + 1]) / scale).max() < 1e-9, (fname, c, t)
            assert abs(rew - R[t]) <= 1e-9 * max(1.0, abs(R[t]))
            assert done == bool(D[t])
            np.testing.assert_allclose(obs, O[t + 1], rtol=1e-8, atol=1e-9)
            assert env.curr_step == t + 1
            if done:
                break


def test_reset_contracts():
    """Pyrado/tests/test_environments.py:192-213: reset obs in obs_space, state in state_space, same init -> same obs;
    batched envs return device tensors."""
    for cls, kw in ((QQubeSwingUpSim, dict(dt=0.01)), (QCartPoleSwingUpSim, dict(dt=0.01)), (QBallBalancerSim, dict(dt=0.01)),
                    (BallOnBeamSim, dict(dt=0.01)), (OneMassOscillatorSim, dict(dt=0.02))):
        pyrado.set_seed(3)
        env = cls(max_steps=50, **kw)
        obs = env.reset()
        assert env.obs_space.contains(obs) and env.state_space.contains(env.state)
        assert env.init_space.contains(env.state[: env.init_space.flat_dim]) or cls is QBallBalancerSim
        init = env.init_space.sample_uniform()
        o1, o2 = env.reset(init_state=init.copy()), env.reset(init_state=init.copy())
        np.testing.assert_array_equal(o1, o2)
        benv = cls(max_steps=50, num_envs=33, **kw)
        bo = benv.reset()
        assert isinstance(bo, torch.Tensor) and bo.is_cuda and bo.shape == (33, env.obs_space.flat_dim)
        assert benv.state.shape == (33, env.state_space.flat_dim) and float(benv.state.std(dim=0).max()) > 0
        o, r, d, info = benv.step(torch.zeros(33, env.act_space.flat_dim, device=bo.device))
        assert o.shape == bo.shape and r.shape == (33,) and d.dtype == torch.bool and int(info["status"].abs().sum()) == 0


def test_nan_action_raises_value_err():
    env = QQubeSwingUpSim(dt=0.01, max_steps=10)
    env.reset()
    with pytest.raises(pyrado.ValueErr):
        env.step(np.array([np.nan]))


def test_wrappers_fused_equal_unfused():
    """ActNormWrapper through the kernel flag == the formula applied on the host (action_normalization.py:69-75)."""
    a = ActNormWrapper(QQubeSwingUpSim(dt=0.004, max_steps=100))
    b = QQubeSwingUpSim(dt=0.004, max_steps=100)
    init = np.array([0.01, -0.01, 0.0, 0.0])
    oa, ob = a.reset(init_state=init), b.reset(init_state=init)
    np.testing.assert_array_equal(oa, ob)
    rng = np.random.RandomState(0)
    for _ in range(30):
        act = rng.uniform(-1.2, 1.2, size=1)
        oa, ra, da, _ = a.step(act)
        ob, rb, db, _ = b.step(4.5 * act)
        np.testing.assert_allclose(oa, ob, rtol=1e-13, atol=1e-13)
        assert ra == pytest.approx(rb, rel=1e-13) and da == db
    # observation noise: zero std leaves the observation alone, a non-zero one moves it by about std
    n0 = GaussianObsNoiseWrapper(QQubeSwingUpSim(dt=0.004, max_steps=100), noise_std=np.zeros(6))
    np.testing.assert_array_equal(n0.reset(init_state=init), ob * 0 + b.reset(init_state=init))
    n1 = GaussianObsNoiseWrapper(QQubeSwingUpSim(dt=0.004, max_steps=100, num_envs=4096), noise_std=0.1 * np.ones(6),
                                 noise_mean=0.5 * np.ones(6))
    clean = QQubeSwingUpSim(dt=0.004, max_steps=100, num_envs=4096)
    diff = (n1.reset(init_state=init) - clean.reset(init_state=init)).double()
    assert abs(float(diff.mean()) - 0.5) < 0.01 and abs(float(diff.std()) - 0.1) < 0.01


def test_domain_randomizer_and_wrappers():
    """Pyrado/tests/test_domain_randomization.py:59-123 and environment_wrappers/test_domain_randomization.py:45-88."""
    pyrado.set_seed(0)
    r = DomainRandomizer(NormalDomainParam(name="mass", mean=1.2, std=0.1, clip_lo=1e-3),
                         UniformDomainParam(name="length", mean=4.0, halfspan=0.5, roundint=True))
    for dp in r.domain_params:
        s = dp.sample(5)
        assert len(s) == 5 and all(isinstance(x, torch.Tensor) and x.ndim == 0 for x in s)
    r.randomize(10)
    lst = r.get_params(-1, fmt="list", dtype="numpy")
    dct = r.get_params(-1, fmt="dict", dtype="torch")
    assert isinstance(lst, list) and len(lst) == 10 and lst[0]["mass"].dtype == np.float32 and lst[0]["length"].dtype == np.int32
    assert isinstance(dct, dict) and len(dct["mass"]) == 10 and isinstance(dct["mass"][0], torch.Tensor)
    assert float(dct["mass"][3]) == float(lst[3]["mass"])
    assert isinstance(r.get_params(1, fmt="dict"), dict) and len(r.get_params(3, fmt="list")) == 3
    assert all(3.0 <= float(x["length"]) <= 5.0 and float(x["length"]) == round(float(x["length"])) for x in lst)
    r.randomize(20000)
    m = np.array([float(x) for x in r.get_params(-1, fmt="dict")["mass"]])
    assert abs(m.mean() - 1.2) < 5e-3 and abs(m.std() - 0.1) < 5e-3
    # Live: new parameters at every reset; Buffer: cycles
    env = QQubeSwingUpSim(dt=0.01, max_steps=20)
    live = DomainRandWrapperLive(env, create_default_randomizer(env))
    seen = []
    for _ in range(3):
        live.reset()
        seen.append(live.domain_param["mass_pend_pole"])
    assert len(set(seen)) == 3 and all(v > 0 for v in seen)
    buf = DomainRandWrapperBuffer(QQubeSwingUpSim(dt=0.01, max_steps=20), create_default_randomizer(env))
    buf.fill_buffer(3)
    vals = []
    for i in range(6):
        assert buf.ring_idx == i % 3
        buf.reset()
        vals.append(buf.domain_param["mass_pend_pole"])
    assert vals[:3] == vals[3:] and len(set(vals[:3])) == 3
    assert vals[0] == pytest.approx(float(buf.buffer[0]["mass_pend_pole"]))
    # batch: one parameter set per env, written on the device
    benv = QQubeSwingUpSim(dt=0.01, max_steps=20, num_envs=1000)
    blive = DomainRandWrapperLive(benv, create_default_randomizer(benv))
    blive.reset()
    mp = blive.domain_param["mass_pend_pole"]
    assert mp.shape == (1000,) and abs(mp.mean() - 0.024) < 1e-3 and abs(mp.std() - 0.0048) < 5e-4
    assert np.all(blive.domain_param["voltage_thold_neg"] == 0.0)  # not randomized: nominal everywhere


class _Transparent(EnvWrapperObs):
    """An observation wrapper the kernels do not know: forces rollout() onto the generic stepping path."""

    def _process_obs(self, obs):
        return obs


def _chain(num_envs=1, noise=False, dr=True, max_steps=80):
    env = QQubeSwingUpSim(dt=1 / 250.0, max_steps=max_steps, num_envs=num_envs)
    out = ActNormWrapper(env)
    if noise:
        out = GaussianObsNoiseWrapper(out, noise_std=0.01 * np.ones(6))
    if dr:
        out = DomainRandWrapperLive(out, create_default_randomizer(env))
    return out


def _policy(env, hidden=16):
    torch.manual_seed(1)
    return FNNPolicy(env.spec, hidden_sizes=[hidden, hidden], hidden_nonlin=torch.tanh)


def test_rollout_returns_reference_shaped_step_sequence():
    env = _chain()
    pol = _policy(env)
    ro = rollout(env, pol, eval=True, seed=0, sub_seed=0, sub_sub_seed=3)
    assert isinstance(ro, StepSequence) and 1 <= len(ro) <= 80 and ro.done[-1] and not ro.done[:-1].any()
    T = len(ro)
    assert ro.observations.shape == (T + 1, 6) and ro.states.shape == (T + 1, 4) and ro.actions.shape == (T, 1)
    assert ro.actions_applied.shape == (T, 1) and ro.rewards.shape == (T,) and ro.time.shape == (T + 1,)
    assert ro.time[1] == pytest.approx(0.004) and np.all(np.abs(ro.actions_applied) <= 1.0)
    assert ro.rollout_info["env_name"] == "qq-su" and set(ro.rollout_info["domain_param"]) >= set(env.supported_domain_param)
    assert ro.rollout_info["domain_param"]["mass_pend_pole"] != 0.024  # randomized
    assert env.curr_step == T and np.allclose(env.state, ro.states[-1])
    # observation = observe(state) (no noise wrapper here)
    np.testing.assert_allclose(ro.observations[:, 0], np.sin(ro.states[:, 0]), atol=1e-12)
    # the reward is the task's reward of (state, unclipped de-normalised action) -- recompute two rows on the device
    same = rollout(env, pol, eval=True, seed=0, sub_seed=0, sub_sub_seed=3)
    np.testing.assert_array_equal(same.rewards, ro.rewards)
    other = rollout(env, pol, eval=True, seed=0, sub_seed=0, sub_sub_seed=4)
    assert not np.array_equal(other.states[0], ro.states[0])  # test_sampling.py:625-646: rollouts differ by index


@pytest.mark.parametrize("noise", [False, True])
def test_fused_rollout_equals_stepping_loop(noise):
    """One persistent-kernel launch == the reference's loop over env.step (same Philox keys for DR, init state and
    observation noise); fp64 env, the fp32 policy is evaluated by torch on the host in the generic path."""
    pol = None
    ros = []
    for generic in (False, True):
        env = _chain(noise=noise)
        pol = pol or _policy(env)
        if generic:
            env = _Transparent(env)
        ros.append(rollout(env, pol, eval=True, seed=5, sub_seed=1, sub_sub_seed=2))
    a, b = ros
    assert a.rollout_info["domain_param"]["gravity_const"] == pytest.approx(b.rollout_info["domain_param"]["gravity_const"], rel=1e-7)
    np.testing.assert_allclose(a.states[0], b.states[0], rtol=0, atol=1e-15)
    n = min(len(a), len(b), 40)
    np.testing.assert_allclose(a.states[:n], b.states[:n], rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(a.observations[:n], b.observations[:n], rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(a.rewards[:n], b.rewards[:n], rtol=1e-4, atol=1e-5)
    # with an idle policy the two paths agree to the last bit over the whole episode
    idle = []
    for generic in (False, True):
        env = _chain(noise=noise)
        e2 = _Transparent(env) if generic else env
        idle.append(rollout(e2, IdlePolicy(env.spec), eval=True, seed=5, sub_seed=1, sub_sub_seed=2))
    assert len(idle[0]) == len(idle[1])
    np.testing.assert_allclose(idle[0].states, idle[1].states, rtol=1e-13, atol=1e-13)
    np.testing.assert_allclose(idle[0].rewards, idle[1].rewards, rtol=1e-13, atol=1e-13)


def test_sampler_sequential_equals_parallel_and_split_invariance():
    """Pyrado/tests/test_sampling.py:570-701 restated: rollout(seed, sub_seed=0, sub_sub_seed=i) equals rollout i of
    the sampler; the result is identical for any lane chunking and any number of shards (GPUs)."""
    env = _chain()
    pol = _policy(env)
    n = 24
    full = ParallelRolloutSampler(env, pol, num_workers=4, min_rollouts=n, seed=0).sample(eval=True)
    assert len(full) == n and all(isinstance(r, StepSequence) for r in full)
    seq = [rollout(env, pol, eval=True, seed=0, sub_seed=0, sub_sub_seed=i) for i in range(n)]
    chunked = ParallelRolloutSampler(env, pol, num_workers=1, min_rollouts=n, seed=0, max_lanes=7).sample(eval=True)
    sharded = []
    for r in range(3):
        sharded += ParallelRolloutSampler(env, pol, 1, min_rollouts=n, seed=0, shard=(r, 3)).sample(eval=True)
    for other in (seq, chunked, sharded):
        assert len(other) == n
        for a, b in zip(full, other):
            assert len(a) == len(b)
            np.testing.assert_array_equal(a.states, b.states)
            np.testing.assert_array_equal(a.rewards, b.rewards)
            np.testing.assert_array_equal(a.actions, b.actions)
            assert a.rollout_info["domain_param"] == b.rollout_info["domain_param"]
    assert len({r.states[0].tobytes() for r in full}) == n  # rollouts within one batch differ
    # the next sample() call uses sub_seed 1: different rollouts
    s = ParallelRolloutSampler(env, pol, 1, min_rollouts=4, seed=0)
    first, second = s.sample(eval=True), s.sample(eval=True)
    assert not np.array_equal(first[0].states[0], second[0].states[0])
    st = s.statistics()
    assert st["num_rollouts"] == 4 and st["num_steps"] == sum(len(r) for r in second)
    assert st["avg_return"] == pytest.approx(np.mean([r.undiscounted_return() for r in second]), rel=1e-9)


def test_sampler_min_steps_init_states_and_domain_params():
    env = _chain(dr=False, max_steps=50)
    pol = _policy(env)
    s = ParallelRolloutSampler(env, pol, 2, min_steps=333, seed=1)
    ros = s.sample(eval=True)
    total = sum(len(r) for r in ros)
    assert total >= 333 and total - len(ros[-1]) < 333  # cut right after the rollout that reaches the minimum
    again = ParallelRolloutSampler(env, pol, 2, min_steps=333, seed=1, max_lanes=3).sample(eval=True)
    assert [len(r) for r in again] == [len(r) for r in ros]
    np.testing.assert_array_equal(again[-1].states, ros[-1].states)
    # explicit init states x domain params: every combination, repeated up to min_rollouts (:285-303)
    inits = [np.array([0.01, 0.0, 0.0, 0.0]), np.array([-0.02, 0.01, 0.0, 0.0])]
    dps = [dict(mass_pend_pole=0.02), dict(mass_pend_pole=0.03), dict(mass_pend_pole=0.04)]
    ros = ParallelRolloutSampler(env, pol, 2, min_rollouts=7, seed=1).sample(init_states=inits, domain_params=dps, eval=True)
    assert len(ros) == 12
    for i, ro in enumerate(ros):
        np.testing.assert_allclose(ro.states[0], inits[(i % 6) // 3])
        assert ro.rollout_info["domain_param"]["mass_pend_pole"] == pytest.approx(dps[i % 3]["mass_pend_pole"])
    np.testing.assert_array_equal(ros[0].states, ros[6].states)  # same reset arguments, deterministic env
    with pytest.raises(NotImplementedError):
        s.sample(init_states=inits)


def test_sampler_linear_policy_and_generic_fallback():
    env = _chain(dr=False, max_steps=30)
    lin = LinearPolicy(env.spec, FeatureStack(const_feat, identity_feat, squared_feat))
    a = ParallelRolloutSampler(env, lin, 1, min_rollouts=5, seed=2).sample(eval=True)
    b = ParallelRolloutSampler(_Transparent(env), lin, 1, min_rollouts=5, seed=2).sample(eval=True)  # generic path
    assert len(a) == len(b) == 5
    for x, y in zip(a, b):
        np.testing.assert_allclose(x.states[0], y.states[0], atol=1e-15)
        m = min(len(x), len(y), 15)
        np.testing.assert_allclose(x.states[:m], y.states[:m], rtol=1e-4, atol=1e-5)


def test_sample_batch_device_resident():
    env = _chain(num_envs=1, max_steps=120)
    pol = _policy(env, hidden=64)
    inner = env.wrapped_env.wrapped_env
    inner._dtype = torch.float32  # production build
    s = ParallelRolloutSampler(env, pol, 1, min_rollouts=1, seed=0)
    res = s.sample_batch(5000, auto_reset=True)
    assert res["ret"].is_cuda and res["length"].shape == (5000,)
    st = s.statistics()
    # every closing row of an episode that ended early takes one of the 121 rows of its lane
    assert 0.98 * 5000 * 120 <= st["num_steps"] <= 5000 * 120 and st["num_rollouts"] >= 5000
    assert st["num_steps"] + st["num_rollouts"] >= 5000 * 120
    assert st["min_return"] <= st["avg_return"] <= st["max_return"]
    ret_h, len_h, st_h = s.read_results(res)  # pinned staging buffers of the sampler
    np.testing.assert_array_equal(ret_h, res["ret"].cpu().numpy())
    np.testing.assert_array_equal(len_h, res["length"].cpu().numpy())
    np.testing.assert_array_equal(st_h, s.last_stats.cpu().numpy())
    rec = s.sample_batch(512, auto_reset=False, record=True)
    assert rec["traj"].shape[:2] == (121, 512) and int(rec["length"].max()) <= 120
    assert s.read_results(rec)[0].shape == (512,)


def test_step_diff_state_autograd_matches_reference():
    """sac_wgrad's call pattern (sac_wgrad/main.py:296-298): ObsActCat(ActNorm(QCartPole)).step_diff_state in a loop,
    torch.autograd through the adjoint kernel; gradients vs autograd on the unmodified reference (golden)."""
    g = golden("step_diff_qcp")
    key = "full_f64"
    s0, acts = g[f"{key}_s0"], g[f"{key}_acts"]
    H = acts.shape[0]
    env = ObsActCatWrapper(ActNormWrapper(QCartPoleSwingUpSim(dt=1 / 20.0, max_steps=300, long=True, wild_init=True)))
    dev = env.wrapped_env.wrapped_env.device
    state = torch.tensor(s0, dtype=torch.float64, device=dev, requires_grad=True)
    a_all = torch.tensor(acts, dtype=torch.float64, device=dev, requires_grad=True)
    s, thdd, total = state, None, 0.0
    for t in range(H):
        s, rew, done, info = env.step_diff_state(s, a_all[t], thdd)
        thdd = info["th_ddot"]
        total = total + rew.sum()
        np.testing.assert_allclose(s.detach().cpu().numpy(), g[f"{key}_states"][t], rtol=1e-9, atol=1e-9)
        assert done.dtype == torch.bool
    (total + 0.1 * s.sum()).backward()
    ref_s0, ref_a = g[f"{key}_g_s0"], g[f"{key}_g_a"]
    assert np.abs(state.grad.cpu().numpy() - ref_s0).max() < 1e-8 * max(np.abs(ref_s0).max(), 1.0)
    assert np.abs(a_all.grad.cpu().numpy() - ref_a).max() < 1e-8 * max(np.abs(ref_a).max(), 1.0)
