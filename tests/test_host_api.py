"""
CPU tests of the host-side mirror of the reference API (no CUDA device needed, no compute calls): seed hashing,
spaces, task descriptors, StepSequence, wrapper-chain logic, randomizer tables, sharding arithmetic, and the two-process
gloo test of the statistics all-reduce / ragged all-gather that the multi-GPU sampler uses.
Reference known answers: Pyrado/tests/test_set_seed.py:34-55, Pyrado/tests/test_tasks.py:82-99 (in test_oracle_golden).
"""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

import simurlacra_b200 as pyrado
from conftest import ROOT, golden
from simurlacra_b200 import _lib as L
from simurlacra_b200.domain_randomization import (BernoulliDomainParam, DomainRandomizer, NormalDomainParam,
                                                   UniformDomainParam)
from simurlacra_b200.domain_randomization.default_randomizers import (create_default_randomizer,
                                                                       create_default_randomizer_qbb,
                                                                       create_default_randomizer_qq,
                                                                       create_zero_var_randomizer)
from simurlacra_b200.environment_wrappers import (ActNormWrapper, DomainRandWrapperBuffer, DomainRandWrapperLive,
                                                   EnvWrapperAct, GaussianObsNoiseWrapper, inner_env, remove_env,
                                                   typed_env)
from simurlacra_b200.environments.pysim.ball_on_beam import BallOnBeamSim
from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
from simurlacra_b200.environments.pysim.quanser_cartpole import QCartPoleSwingUpSim
from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
from simurlacra_b200.policies import FeatureStack, FNNPolicy, IdlePolicy, LinearPolicy
from simurlacra_b200.policies.features import const_feat, identity_feat, squared_feat
from simurlacra_b200.sampling import sharding
from simurlacra_b200.sampling.rollout import compile_chain
from simurlacra_b200.sampling.step_sequence import StepSequence, discounted_reverse_cumsum, gae_returns

_SEED_KA = {0: [813134492, 4276188331, 229607210, 3918913762], 1: [532102107, 2754337450, 713485941, 3511146676]}
# all 18 vectors of Pyrado/tests/test_set_seed.py:34-55: per base seed the answers for (sub, subsub) hashing to
# "b-0-0", "b-0-1", "b-1-0", "b-1-1"; None counts as 0
SEED_VECTORS = [((b, s, ss), _SEED_KA[b][2 * int(bool(s)) + int(bool(ss))])
                for b in (0, 1) for s in (None, 0, 1) for ss in (None, 0, 1)]


@pytest.mark.parametrize("triple,expected", SEED_VECTORS)
def test_seed_hash_known_answers(triple, expected):
    assert pyrado.seed_hash(*triple) == expected
    assert pyrado.set_seed(*triple) == expected
    assert pyrado.device_seed() == expected and pyrado.get_base_seed() == triple[0]


def test_env_ctor_errors_and_spaces():
    with pytest.raises(pyrado.ValueErr):
        QQubeSwingUpSim(dt=-0.1)
    with pytest.raises(pyrado.TypeErr):
        QQubeSwingUpSim(dt="a")
    with pytest.raises(pyrado.ValueErr):
        QQubeSwingUpSim(dt=0.01, max_steps=0)
    env = QQubeSwingUpSim(dt=1 / 250.0, max_steps=600)
    assert env.obs_space.flat_dim == 6 and env.act_space.flat_dim == 1 and env.state_space.flat_dim == 4
    assert env.spec.obs_space is env.obs_space
    assert set(env.supported_domain_param) == set(env.get_nominal_domain_param())
    for _ in range(100):  # Pyrado/tests/test_environments.py:140-148
        assert env.init_space.contains(env.init_space.sample_uniform())
    env.max_steps = 10
    assert env.max_steps == 10
    with pytest.raises(pyrado.ValueErr):
        env.max_steps = 0


def test_domain_param_set_get_roundtrip_and_dependent_spaces():
    """Pyrado/tests/test_domain_randomization.py:194-205: set every key, read it back."""
    for env in (QQubeSwingUpSim(dt=0.01), QCartPoleSwingUpSim(dt=0.01, long=True), QBallBalancerSim(dt=0.01),
                BallOnBeamSim(dt=0.01)):
        nom = env.get_nominal_domain_param() if not isinstance(env, QCartPoleSwingUpSim) else env.domain_param
        for k, v in nom.items():
            new = v * 1.1 + 0.01
            env.domain_param = {k: new}
            assert abs(env.domain_param[k] - new) < 5e-4
    cp = QCartPoleSwingUpSim(dt=0.01)
    cp.domain_param = dict(rail_length=1.0)
    assert cp.state_space.bound_up[0] == pytest.approx(0.5 - 0.15)
    bob = BallOnBeamSim(dt=0.01)
    bob.domain_param = dict(beam_length=4.0)
    assert bob.state_space.bound_up[0] == pytest.approx(2.0)


def test_wild_init_string_semantics():
    wide = QCartPoleSwingUpSim(dt=0.01, wild_init="True").init_space.bound_up
    narrow = QCartPoleSwingUpSim(dt=0.01, wild_init="False").init_space.bound_up
    other = QCartPoleSwingUpSim(dt=0.01, wild_init=True).init_space.bound_up  # a bool takes the else branch
    assert wide[1] == pytest.approx(np.pi) and wide[0] == pytest.approx(0.25)
    assert narrow[1] == pytest.approx(2 / 180 * np.pi)
    assert other[0] == pytest.approx(0.02) and other[1] == pytest.approx(np.pi)


def test_domain_param_classes_and_randomizer_table():
    with pytest.raises(ValueError):
        NormalDomainParam(name="m", mean=1.0, std=0.0)
    with pytest.raises(ValueError):
        UniformDomainParam(name="m", mean=1.0, halfspan=0.0)
    with pytest.raises(ValueError):
        BernoulliDomainParam(name="b", val_0=0, val_1=1, prob_1=1.5)
    b = BernoulliDomainParam(name="b", val_0=2.0, val_1=4.0, prob_1=0.25)
    assert b.mean == pytest.approx(2.5)
    n = NormalDomainParam(name="m", mean=1.0, std=0.2, clip_lo=1e-3)
    n.adapt("std", 0.1)
    assert n.std == 0.1
    with pytest.raises(pyrado.KeyErr):
        n.adapt("halfspan", 0.1)
    r = create_default_randomizer_qq()
    assert len(r.domain_params) == 9 and "gravity_const" in str(r)
    env = QQubeSwingUpSim(dt=0.01)
    tab = r.table(env._param_names)
    assert tab._n == 9
    assert [tab[i].param_index for i in range(9)] == [env._param_names.index(dp.name) for dp in r.domain_params]
    assert tab[0].kind == L.DISTR_NORMAL and tab[0].a == pytest.approx(9.81) and tab[0].b == pytest.approx(0.981)
    assert tab[0].clip_lo == pytest.approx(1e-3)
    q = create_default_randomizer_qbb()
    assert len(q.domain_params) == 20
    kinds = [dp._distr_kind for dp in q.domain_params]
    assert kinds.count("normal") == 10 and kinds.count("uniform") == 10
    r.rescale_distr_param("std", 0.5)
    assert r.domain_params[0].std == pytest.approx(0.4905)
    z = create_zero_var_randomizer(ActNormWrapper(env))
    assert z.domain_params[0].std == pytest.approx(0.981e-4)
    assert type(create_default_randomizer(QCartPoleSwingUpSim(dt=0.01))).__name__ == "DomainRandomizer"
    with pytest.raises(pyrado.TypeErr):
        DomainRandomizer("not a param")


def test_wrapper_chain_delegation_and_fusion():
    env = QQubeSwingUpSim(dt=1 / 250.0, max_steps=600)
    an = ActNormWrapper(env)
    assert np.all(an.act_space.bound_up == 1) and np.all(an.act_space.bound_lo == -1)
    assert an._process_act(np.array([0.5]))[0] == pytest.approx(2.25)  # lb + (a + 1)(ub - lb)/2 with +-4.5 V
    assert float(an._process_act_tensor(torch.tensor([[1.0]]))[0, 0]) == pytest.approx(4.5)
    noisy = GaussianObsNoiseWrapper(an, noise_std=0.01 * np.ones(6))
    with pytest.raises(pyrado.ShapeErr):
        GaussianObsNoiseWrapper(an, noise_std=np.ones(3))
    dr = DomainRandWrapperLive(noisy, create_default_randomizer_qq())
    assert inner_env(dr) is env and typed_env(dr, ActNormWrapper) is an and dr.name == "qq-su"
    assert dr.max_steps == 600 and dr.dt == pytest.approx(0.004)
    dp = dr.domain_param
    assert "obs_noise_std" in dp and dp["gravity_const"] == pytest.approx(9.81)
    dr.domain_param = dict(obs_noise_std=0.5 * np.ones(6), gravity_const=9.0)
    assert noisy._std[0] == 0.5 and env.domain_param["gravity_const"] == 9.0
    cc = compile_chain(dr)
    assert cc is not None and cc.sim is env and cc.fusion.act_norm and cc.fusion.noise_std[0] == 0.5
    assert cc.live_randomizer is dr.randomizer
    cfg = env._make_cfg(cc.fusion)
    assert cfg.act_norm == 1 and cfg.obs_noise == 1 and cfg.noise_std[5] == 0.5 and cfg.kind == L.QQ
    stripped = remove_env(dr, GaussianObsNoiseWrapper)
    assert typed_env(stripped, GaussianObsNoiseWrapper) is None and inner_env(stripped) is env
    assert typed_env(dr, GaussianObsNoiseWrapper) is noisy  # the original chain is untouched

    class Scale(EnvWrapperAct):  # a user wrapper the kernels do not know
        def _process_act(self, act):
            return 2 * act

    assert compile_chain(Scale(an)) is None                      # not fusable -> generic stepping path
    assert compile_chain(ActNormWrapper(Scale(env))) is None
    with pytest.raises(pyrado.TypeErr):
        DomainRandWrapperLive(env, "nope")
    buf = DomainRandWrapperBuffer(env, None)
    buf.buffer = [dict(gravity_const=1.0), dict(gravity_const=2.0)]
    buf._ring_idx = 0
    assert buf._next_set()["gravity_const"] == 1.0 and buf._next_set()["gravity_const"] == 2.0 and buf.ring_idx == 0
    with pytest.raises(pyrado.ValueErr):
        DomainRandWrapperBuffer(env, None, selection="zigzag")


def test_policies_descriptors_and_torch_forward():
    env = QQubeSwingUpSim(dt=0.01)
    torch.manual_seed(0)
    pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
    assert pol.num_param == 6 * 16 + 16 + 16 * 16 + 16 + 16 + 1
    assert pol(torch.zeros(6)).shape == (1,) and pol(torch.zeros(3, 6)).shape == (3, 1)  # test_policies.py:361-455
    v = pol.param_values.clone()
    pol.param_values = v * 0 + 0.5
    assert torch.all(pol.param_values == 0.5)
    with pytest.raises(pyrado.ShapeErr):
        pol.param_values = torch.zeros(3)
    assert pol.net.engine_codes() == (0, 0)
    assert FNNPolicy(env.spec, [8], torch.relu, output_nonlin=torch.tanh).net.engine_codes() == (1, 1)
    assert FNNPolicy(env.spec, [8], torch.sigmoid).net.engine_codes() is None  # no kernel -> generic path
    fork = FNNPolicy(QCartPoleSwingUpSim(dt=0.01).spec._replace(obs_space=QCartPoleSwingUpSim(dt=0.01).state_space),
                     [16], torch.tanh, cartpole_feats=True)
    assert fork.net.sizes[0] == 5 and fork(torch.zeros(2, 4)).shape == (2, 1)
    lin = LinearPolicy(env.spec, FeatureStack(const_feat, identity_feat, squared_feat))
    assert lin.num_active_feat == 13 and lin.features.engine_codes() == ["const", "identity", "squared"]
    assert lin(torch.ones(6)).shape == (1,) and lin(torch.ones(4, 6)).shape == (4, 1)
    assert FeatureStack(identity_feat, lambda x: x).engine_codes() is None
    assert torch.all(IdlePolicy(env.spec)(torch.ones(5, 6)) == 0)
    with pytest.raises(pyrado.TypeErr):
        LinearPolicy(env.spec, "feats")


def test_step_sequence_container():
    rng = np.random.RandomState(0)
    ro = StepSequence(rewards=rng.rand(5), observations=rng.randn(6, 3), actions=rng.randn(5, 1), states=rng.randn(6, 2),
                      rollout_info=dict(env_name="x"))
    assert len(ro) == 5 and ro.done[-1] and list(ro.rollout_bounds) == [0, 5] and ro.rollout_count == 1
    assert ro[2].reward == ro.rewards[2] and np.all(ro[2].next_observation == ro.observations[3]) and not ro[2].done
    assert ro[-1].done and np.all(ro[1:3].observations == ro.observations[1:3])
    with pytest.raises(pyrado.ShapeErr):
        StepSequence(rewards=[], observations=[], actions=[])
    with pytest.raises(pyrado.ShapeErr):
        ro.add_data("bad", np.zeros(3))
    with pytest.raises(pyrado.KeyErr):
        ro.add_data("states", np.zeros(5))
    ro.add_data("values", rng.rand(6))
    c = StepSequence.concat([ro, ro])
    assert c.length == 10 and list(c.rollout_bounds) == [0, 5, 10] and c.rollout_info == dict(env_name="x")
    assert np.all(c.get_rollout(1).rewards == ro.rewards) and len(list(c.iterate_rollouts())) == 2
    assert c.observations.shape == (10, 3)  # the final observation of every part is dropped
    assert ro.undiscounted_return() == pytest.approx(ro.rewards.sum())
    assert ro.discounted_return(0.9) == pytest.approx(np.dot(ro.rewards, 0.9 ** np.arange(5)))
    y = discounted_reverse_cumsum(ro.rewards, 0.9)
    assert y[0] == pytest.approx(ro.discounted_return(0.9)) and y[-1] == pytest.approx(ro.rewards[-1])
    adv = gae_returns(ro, gamma=0.99, lamb=1.0)  # lambda = 1: discounted return minus the baseline (+ bootstrap)
    assert adv[-1] == pytest.approx(ro.rewards[-1] - ro.values[4])
    sizes = [b.length for b in c.split_ordered_batches(batch_size=4)]
    assert sizes == [4, 4, 2] and sum(b.length for b in c.split_shuffled_batches(3)) == 10
    assert all(b.rollout_count >= 1 for b in c.split_shuffled_batches(4, complete_rollouts=True))
    StepSequence.pad(ro, 8)
    assert ro.length == 8 and ro.observations.shape == (9, 3) and list(ro.rollout_bounds) == [0, 8] and not ro.done[-1]
    ro.torch()
    assert isinstance(ro.rewards, torch.Tensor) and ro.data_format == "torch"
    ro.numpy()
    assert isinstance(ro.rewards, np.ndarray)


def test_shard_ranges_cover_and_balance():
    for n in (0, 1, 7, 8, 1000, 1 << 20):
        for ws in (1, 2, 3, 4, 8):
            blocks = [sharding.shard_range(n, r, ws) for r in range(ws)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(ws - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    s = torch.tensor([4, 10.0, 30.0, 40, 40, 1, 1.0, 4.0], dtype=torch.float64)
    d = sharding.summarize(s)
    assert d["avg_return"] == 2.5 and d["std_return"] == pytest.approx((7.5 - 6.25) ** 0.5) and d["avg_rollout_len"] == 10


_GLOO_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.environ["SRE_ROOT"])
from simurlacra_b200.sampling import sharding
dist.init_process_group("gloo")
rank, ws = sharding.world()
n_total = 7
lo, hi = sharding.shard_range(n_total, rank, ws)
ret = torch.arange(lo, hi, dtype=torch.float64) + 1.0          # "returns" of this rank's rollouts
stats = torch.tensor([hi - lo, ret.sum(), (ret ** 2).sum(), 10.0 * (hi - lo), 10.0 * (hi - lo), 0.0, ret.min(), ret.max()],
                     dtype=torch.float64)
sharding.allreduce_stats(stats)
assert stats.tolist() == [7.0, 28.0, 140.0, 70.0, 70.0, 0.0, 1.0, 7.0], stats
full = sharding.allgather_ragged(ret, n_total)
assert full.tolist() == [1.0, 2.0, 3.0, 4.0, 5.0, 6.0, 7.0], full
T, W = 3, 2                                                     # record shards -> the learner (rank 0)
traj = (torch.arange(lo, hi, dtype=torch.float32).view(1, -1, 1) + 100.0 * torch.arange(T).view(T, 1, 1)).expand(T, hi - lo, W).contiguous()
flag = torch.arange(lo, hi, dtype=torch.uint8).view(1, -1).expand(T, hi - lo).contiguous()
gt, gf = sharding.gather_records(traj, flag, n_total, dst=0)
if rank == 0:
    assert gt.shape == (T, n_total, W) and gt[2, :, 1].tolist() == [200.0 + i for i in range(n_total)], gt
    assert gf[1].tolist() == list(range(n_total)), gf
else:
    assert gt is None and gf is None
lin = torch.nn.Linear(3, 2)
with torch.no_grad():
    lin.weight.fill_(float(rank + 1))
sharding.broadcast_policy(lin, src=0)
assert float(lin.weight[0, 0]) == 1.0
dist.barrier(); dist.destroy_process_group()
sys.stdout.write(f"rank {rank} ok\n"); sys.stdout.flush()
"""


def test_multi_rank_statistics_allreduce_gloo(tmp_path):
    """world_size 2 on CPU (gloo): the sampler's statistics all-reduce, ragged all-gather and policy broadcast."""
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, SRE_ROOT=ROOT, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout


def test_exploration_strategies_torch_side():
    """Pyrado/tests/test_exploration.py: noise modules and step-based exploration strategies around a policy."""
    from simurlacra_b200.exploration import (DiagNormalNoise, NormalActNoiseExplStrat, UniformActNoiseExplStrat,
                                             UniformNoise)

    env = QBallBalancerSim(dt=0.01)
    torch.manual_seed(0)
    pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
    ex = NormalActNoiseExplStrat(pol, std_init=0.5)
    assert torch.allclose(ex.std, torch.full((2,), 0.5)) and ex.mean is None
    assert ex.noise.engine_noise() == (L.ACT_NOISE_NORMAL, [0.5, 0.5], [0.0, 0.0])
    obs = torch.zeros(8)
    a = ex(obs)
    assert a.shape == (2,) and not torch.equal(a, pol(obs))
    acts = torch.stack([ex(obs) for _ in range(4000)])
    assert abs(float(acts.std(dim=0).mean()) - 0.5) < 0.03 and torch.allclose(acts.mean(dim=0), pol(obs), atol=0.05)
    ex.std = torch.tensor([1e-9, 2.0])  # the setter clips at std_min (normal_noise.py:118-124)
    assert torch.allclose(ex.std, torch.tensor([1e-3, 2.0]))
    ex.reset_expl_params()
    assert torch.allclose(ex.std, torch.full((2,), 0.5))
    ent = ex.get_entropy()
    assert torch.isclose(ent, torch.tensor(0.5 + 0.5 * np.log(2 * np.pi) + 0.5 * np.log(0.5 ** 4), dtype=ent.dtype))
    ex.noise.adapt(std=torch.tensor([0.1, 0.2]))
    assert torch.allclose(ex.std, torch.tensor([0.1, 0.2]))
    with pytest.raises(AssertionError):
        ex.noise.adapt(mean=torch.zeros(2))  # fixed zero mean
    with pytest.raises(pyrado.ShapeErr):
        ex.noise.adapt(std=torch.ones(3))
    with pytest.raises(pyrado.ValueErr):
        DiagNormalNoise(noise_dim=2, std_init=-1.0)
    with pytest.raises(pyrado.TypeErr):
        DiagNormalNoise(noise_dim=2, std_init="0.1")
    with pytest.raises(pyrado.ShapeErr):
        DiagNormalNoise(noise_dim=2, std_init=torch.ones(3))
    tm = NormalActNoiseExplStrat(pol, std_init=0.1, train_mean=True)
    tm.noise.adapt(mean=torch.tensor([0.3, -0.3]))
    assert tm.noise.engine_noise()[2] == pytest.approx([0.3, -0.3])
    assert set(n for n, _ in tm.named_parameters()) >= {"_noise.log_std", "_noise.mean"}
    # the wrapper forwards the engine precision to the wrapped policy
    tm.engine_precision = L.PREC_TF32X3
    assert pol.engine_precision == L.PREC_TF32X3 and tm.engine_precision == L.PREC_TF32X3
    un = UniformActNoiseExplStrat(pol, halfspan_init=0.2)
    acts = torch.stack([un(obs) for _ in range(2000)]) - pol(obs)
    assert float(acts.abs().max()) <= 0.2 + 1e-6 and float(acts.abs().max()) > 0.19
    assert torch.allclose(un.get_entropy(), torch.log(torch.tensor([0.4, 0.4])))
    assert un.noise.engine_noise()[0] == L.ACT_NOISE_UNIFORM
    with pytest.raises(pyrado.ValueErr):
        UniformNoise(False, 2, halfspan_init=0.0)
    # evaluate(): distribution with gradient data for the learner
    ro = StepSequence(observations=torch.zeros(6, 8), actions=torch.zeros(5, 2), rewards=torch.zeros(5), data_format="torch")
    dist = tm.evaluate(ro)
    lp = dist.log_prob(ro.actions).sum()
    lp.backward()
    assert tm.noise.log_std.grad is not None and pol.param_grad.abs().sum() > 0


def test_obs_norm_and_act_noise_wrappers_host_side():
    """Pyrado/tests/test_environment_wrappers.py (ObsNormWrapper, GaussianActNoiseWrapper): bounds handling, spaces,
    domain-parameter entries, and where the wrappers fuse into the kernels."""
    from simurlacra_b200.environment_wrappers import GaussianActNoiseWrapper, ObsNormWrapper

    qq = QQubeSwingUpSim(dt=0.01)
    on = ObsNormWrapper(qq)
    np.testing.assert_array_equal(on.obs_space.bound_lo, -np.ones(6))
    lb, ub = qq.obs_space.bounds
    x = np.array([0.5, -0.5, 1.0, -1.0, 10.0, -62.0])
    np.testing.assert_allclose(on._process_obs(x), (x - lb) / (ub - lb) * 2 - 1)
    t = on._process_obs(torch.as_tensor(x)[None].repeat(3, 1))
    assert t.shape == (3, 6) and torch.allclose(t[0], torch.as_tensor(on._process_obs(x)))
    qcp = QCartPoleSwingUpSim(dt=0.01)
    with pytest.raises(pyrado.ValueErr):
        ObsNormWrapper(qcp)  # unbounded velocities (observation_normalization.py:69-80)
    with pytest.raises(pyrado.ValueErr):
        ObsNormWrapper(qcp, explicit_lb={"x_dot": -5.0}, explicit_ub={"x_dot": 5.0, "theta_dot": 20.0})
    ok = ObsNormWrapper(qcp, explicit_lb={"x_dot": -5.0, "theta_dot": -20.0}, explicit_ub={"x_dot": 5.0, "theta_dot": 20.0})
    assert ok.ov_ub[3] == 5.0 and ok.ov_lb[4] == -20.0 and ok.ov_ub[0] == qcp.obs_space.bound_up[0]
    an = GaussianActNoiseWrapper(ActNormWrapper(qq), noise_std=[0.1])
    assert an._mean.shape == (1,) and an.act_space.bound_up[0] == 1.0
    with pytest.raises(pyrado.ShapeErr):
        GaussianActNoiseWrapper(qq, noise_std=[0.1, 0.2])
    dp = an.domain_param
    assert dp["act_noise_std"][0] == 0.1 and dp["act_noise_mean"][0] == 0.0
    an.domain_param = dict(act_noise_std=np.array([0.3]))
    assert an._std[0] == 0.3
    samples = np.stack([an._process_act(np.zeros(1)) for _ in range(4000)])
    assert abs(samples.std() - 0.3) < 0.02 and abs(samples.mean()) < 0.02
    # fusion: noise outside / inside ActNorm, ObsNorm above the observation noise; everything else takes the generic path
    cc = compile_chain(ObsNormWrapper(GaussianObsNoiseWrapper(an, noise_std=0.1 * np.ones(6))))
    assert cc.fusion.act_norm and cc.fusion.env_act_noise[0] == L.ENV_ACT_NOISE_OUTER and cc.fusion.obs_norm is not None
    cfg = cc.sim._make_cfg(cc.fusion)
    assert cfg.obs_norm == 1 and cfg.env_act_noise == L.ENV_ACT_NOISE_OUTER and cfg.env_act_noise_std[0] == 0.3
    inner = compile_chain(ActNormWrapper(GaussianActNoiseWrapper(QQubeSwingUpSim(dt=0.01), noise_std=[0.2])))
    assert inner.fusion.env_act_noise[0] == L.ENV_ACT_NOISE_INNER and inner.fusion.act_norm
    assert compile_chain(GaussianObsNoiseWrapper(ObsNormWrapper(QQubeSwingUpSim(dt=0.01)), noise_std=np.ones(6))) is None
    assert compile_chain(GaussianActNoiseWrapper(GaussianActNoiseWrapper(QQubeSwingUpSim(dt=0.01), noise_std=[0.2]),
                                                 noise_std=[0.2])) is None


def test_pendulum_env_host_side():
    """PendulumSim (pendulum.py:43-140): constructor checks, parameter-dependent action space, singular init space,
    default randomizer, C-ABI configuration."""
    from simurlacra_b200.environments.pysim.pendulum import PendulumSim

    env = PendulumSim(dt=0.02, max_steps=100, init_state=np.array([0.3, -0.2]))
    assert env.name == "pend" and env.obs_space.flat_dim == 3 and env.state_space.flat_dim == 2
    assert env.act_space.bound_up[0] == 3.5
    np.testing.assert_array_equal(env.init_space.sample_uniform(), [0.3, -0.2])
    env.domain_param = dict(torque_thold=2.0)
    assert env.act_space.bound_up[0] == 2.0 and env.domain_param["pole_mass"] == 1.0
    with pytest.raises(pyrado.ShapeErr):
        PendulumSim(dt=0.02, init_state=np.zeros(3))
    rz = create_default_randomizer(env)
    assert [p.name for p in rz.domain_params] == list(PendulumSim._param_names)
    cfg = env._make_cfg()
    assert cfg.kind == L.PEND and cfg.has_init_box == 1 and cfg.init_box_lo[0] == 0.3 and cfg.init_box_up[1] == -0.2
    assert cfg.task.rew_kind == L.REW_EXP and cfg.task.radial_mask == 0b10 and cfg.task.state_des[0] == np.pi


def test_parameter_exploration_sampler_host_side():
    """ParameterExplorationSampler (parameter_exploration_sampler.py:151-216): constructor checks, the DR wrapper is
    taken off the chain and kept as the source of domains, the result containers, and the grouped policy descriptor."""
    from simurlacra_b200.policies import FNNPolicy
    from simurlacra_b200.sampling import ParameterExplorationSampler, ParameterSample, ParameterSamplingResult

    sim = QQubeSwingUpSim(dt=1 / 250.0, max_steps=50)
    env = DomainRandWrapperLive(ActNormWrapper(sim), create_default_randomizer(sim))
    pol = FNNPolicy(env.spec, hidden_sizes=[8], hidden_nonlin=torch.tanh)
    for bad in (dict(num_init_states_per_domain=1.0, num_domains=2), dict(num_init_states_per_domain=1, num_domains="2")):
        with pytest.raises(pyrado.TypeErr):
            ParameterExplorationSampler(env, pol, **bad)
    for bad in (dict(num_init_states_per_domain=0, num_domains=2), dict(num_init_states_per_domain=1, num_domains=0)):
        with pytest.raises(pyrado.ValueErr):
            ParameterExplorationSampler(env, pol, **bad)
    pes = ParameterExplorationSampler(env, pol, num_init_states_per_domain=2, num_domains=3, seed=0)
    assert pes.num_rollouts_per_param == 6 and isinstance(pes.env, ActNormWrapper) and pes._dr_wrapper is env
    with pytest.raises(pyrado.ShapeErr):
        pes.sample(torch.zeros(4, pol.num_param + 1))
    with pytest.raises(pyrado.TypeErr):
        pes.sample(torch.zeros(4, pol.num_param), init_states=np.zeros((6, 4)))

    class _Ro:
        def __init__(self, r):
            self.r = r

        def undiscounted_return(self):
            return self.r

    res = ParameterSamplingResult([ParameterSample(params=torch.full((3,), float(i)), rollouts=[_Ro(i), _Ro(i + 1.0)])
                                   for i in range(4)])
    assert len(res) == 4 and res.num_rollouts == 8 and res.parameters.shape == (4, 3)
    np.testing.assert_array_equal(res.mean_returns, [0.5, 1.5, 2.5, 3.5])
    assert res[2].mean_undiscounted_return == 2.5 and res[2].num_rollouts == 2
    assert isinstance(res[1:], ParameterSamplingResult) and len(res[1:]) == 3 and len(res.rollouts[0]) == 2
    with pytest.raises(ValueError):  # grouped descriptors take a matrix of parameter sets
        E_make = __import__("simurlacra_b200.engine", fromlist=["make_policy"]).make_policy
        E_make(L.POL_FNN, torch.zeros(10), sizes=[2, 2, 1], group_size=3)


def test_stateful_step_wrappers_host_side():
    """ActDelay / Downsampling / ObsPartial / ObsVelFilt / StateAugmentation (environment_wrappers/*.py): constructor
    checks, reported spaces, wrapper-owned domain parameters, Space.create_mask (spaces/base.py:103-136)."""
    from simurlacra_b200.environment_wrappers import (ActDelayWrapper, DownsamplingWrapper, ObsPartialWrapper,
                                                       ObsVelFiltWrapper, StateAugmentationWrapper)

    sim = QQubeSwingUpSim(dt=1 / 250.0, max_steps=50)
    sp = sim.obs_space
    np.testing.assert_array_equal(sp.create_mask("theta_dot", 0), [1, 0, 0, 0, 1, 0])
    np.testing.assert_array_equal(sp.create_mask([1, "alpha_dot"]), [0, 1, 0, 0, 0, 1])
    with pytest.raises(pyrado.ValueErr):
        sp.create_mask("nope")
    with pytest.raises(pyrado.ValueErr):
        sp.create_mask(1, 1)
    part = ObsPartialWrapper(sim, idcs=["theta_dot", "alpha_dot"])
    assert part.obs_space.flat_dim == 4 and list(part.obs_space.labels) == ["sin_theta", "cos_theta", "sin_alpha", "cos_alpha"]
    assert ObsPartialWrapper(sim, mask=[1, 1, 0, 0, 0, 0], keep_selected=True).obs_space.flat_dim == 2
    with pytest.raises(pyrado.ShapeErr):
        ObsPartialWrapper(sim, mask=[1, 0])
    np.testing.assert_array_equal(part._process_obs(np.arange(6.0)), [0, 1, 2, 3])
    np.testing.assert_array_equal(part._process_obs(torch.arange(12.0).reshape(2, 6)), [[0, 1, 2, 3], [6, 7, 8, 9]])
    dl = ActDelayWrapper(sim, delay=2)
    assert dl.delay == 2 and dl.domain_param["act_delay"] == 2
    with pytest.raises(pyrado.ValueErr):
        dl.delay = -1
    dl._act_queue = [None, None]
    outs = [dl._process_act(np.array([float(i)])) for i in range(1, 5)]
    np.testing.assert_array_equal(np.concatenate(outs), [0, 0, 1, 2])
    ds = DownsamplingWrapper(sim, factor=3)
    assert ds.domain_param["downsampling"] == 3
    np.testing.assert_array_equal(np.concatenate([ds._process_act(np.array([float(i)])) for i in range(7)]), [0, 0, 0, 3, 3, 3, 6])
    np.testing.assert_allclose([ds._process_obs(np.array([float(i)]))[0] for i in range(5)], [0, 0.5, 1, 2, 3])
    ds.factor = 2
    assert ds._cnt == 0
    vf = ObsVelFiltWrapper(sim, idcs_pos=["theta", "alpha"], idcs_vel=["theta_dot", "alpha_dot"])
    assert vf.b.dtype == np.float32 and vf.z.shape == (1, 2) and vf.obs_space == sim.obs_space
    with pytest.raises(pyrado.ValueErr):
        ObsVelFiltWrapper(sim, idcs_pos=["theta"])
    with pytest.raises(pyrado.ShapeErr):
        ObsVelFiltWrapper(sim, mask_pos=[1, 0], idcs_vel=["theta_dot"])
    # constant position -> zero velocity once the filter starts in steady state; a ramp -> its slope
    vf.init_filter(np.array([0.5, -0.5, 0.0, 0.0]))
    np.testing.assert_allclose(vf._filter(np.array([0.5, -0.5])), 0.0, atol=1e-6)
    from scipy import signal

    vf.init_filter(np.zeros(4))
    x = np.random.RandomState(0).randn(50, 2)
    ys = np.array([vf._filter(xi) for xi in x])
    np.testing.assert_allclose(ys, signal.lfilter(vf.b, vf.a, x, axis=0), rtol=1e-10, atol=1e-10)  # what the reference runs
    aug = StateAugmentationWrapper(sim, domain_param=["mass_pend_pole", "gravity_const"])
    assert aug.obs_space.flat_dim == 8 and list(aug.obs_space.labels[-2:]) == ["mass_pend_pole", "gravity_const"]
    np.testing.assert_allclose(aug.obs_space.bound_up[-2:], 1.5 * aug.nominal)
    assert aug.offset == 6 and aug.mask.sum() == 2
    assert StateAugmentationWrapper(sim).obs_space.flat_dim == 6 + len(sim.domain_param)


def test_discrete_space_and_ball_on_beam_disc():
    """DiscreteSpace (spaces/discrete.py:39-175) and BallOnBeamDiscSim (ball_on_beam.py:139-162)."""
    from simurlacra_b200.environments.pysim.ball_on_beam import BallOnBeamDiscSim
    from simurlacra_b200.spaces import DiscreteSpace

    sp = DiscreteSpace(np.array([[-2.0], [0.0], [2.0]]), labels=["tau"])
    assert sp.num_ele == 3 and sp.flat_dim == 1 and sp.shape == (1,) and sp.bound_abs_up[0] == 2.0
    assert sp.contains(np.array([2.0])) and not sp.contains(np.array([1.0]))
    np.testing.assert_array_equal(sp.project_to(np.array([0.9])), [0.0])
    np.testing.assert_array_equal(sp.project_to(np.array([1.1])), [2.0])
    np.testing.assert_array_equal(sp.project_to(np.array([-7.0])), [-2.0])
    np.testing.assert_array_equal(sp.project_to(np.array([1.0])), [0.0])  # tie -> first of the closest (np.argmin)
    assert sp.project_to(np.array([1e-9]))[0] == 1e-9  # np.isclose to an element: passes unchanged
    with pytest.raises(pyrado.ShapeErr):
        sp.contains(np.zeros(2))
    with pytest.raises(pyrado.ValueErr):
        sp.contains(np.array([np.nan]))
    assert sp.sample_uniform()[0] in (-2.0, 0.0, 2.0)
    assert DiscreteSpace.cat([sp, None, sp]).num_ele == 6 and DiscreteSpace([1, 2, 3]).eles.dtype == np.intc
    env = BallOnBeamDiscSim(dt=0.01, max_steps=100)
    assert env.name == "bob-d" and isinstance(env.act_space, DiscreteSpace)
    np.testing.assert_allclose(env.act_space.eles.reshape(-1), [-29.43, 0.0, 29.43])
    assert env._make_cfg().flags & L.F_DISC_ACT
    env.domain_param = dict(beam_length=1.0)
    np.testing.assert_allclose(env.act_space.eles.reshape(-1), [-14.715, 0.0, 14.715])


def test_discounted_values_return_shapes_match_reference():
    """discounted_values (sampling/step_sequence.py:1075-1090 of the reference): 'torch' -> one 1-D tensor over all
    rollouts, 'numpy' -> one row per rollout; values generated by the unmodified reference (oracle/gen_golden.py)."""
    from simurlacra_b200.sampling.step_sequence import StepSequence, discounted_values

    g = golden("returns_reference")
    for tag in ("eq", "ragged"):
        lens = [int(x) for x in g[f"dvs_{tag}_lens"]]
        rew = g[f"dvs_{tag}_rew"]
        ros, o = [], 0
        for n in lens:
            ros.append(StepSequence(rewards=rew[o:o + n], observations=np.zeros((n + 1, 2)), actions=np.zeros((n, 1)), complete=True))
            o += n
        t = discounted_values(ros, 0.97, "torch")
        assert isinstance(t, torch.Tensor) and t.dim() == 1 and t.shape[0] == sum(lens)
        np.testing.assert_allclose(t.double().numpy(), g[f"dvs_{tag}_torch"], rtol=1e-6, atol=1e-6)
        a = discounted_values(ros, 0.97, "numpy")
        assert len(a) == len(lens)
        if tag == "eq":
            assert a.shape == g["dvs_eq_numpy"].shape
            np.testing.assert_allclose(a, g["dvs_eq_numpy"], rtol=1e-12)


def test_parametrised_features_match_reference_values():
    """MultFeat / ATan2Feat / RFFeat / RBFFeat / sig / bell (policies/features.py:157-161, :181-445) against feature values
    of the unmodified reference (tests/golden/policy_forward.npz, `oracle/gen_golden.py policies`), single and batched
    inputs, the kernel codes they hand to the Linear-policy kernel, and the feature count."""
    import functools

    from conftest import golden
    from simurlacra_b200.policies.features import (ATan2Feat, FeatureStack, MultFeat, RBFFeat, RFFeat, bell_feat, const_feat,
                                                   identity_feat, sig_feat)

    g = golden("policy_forward")
    x = torch.from_numpy(g["pf_x"])
    rf = RFFeat(inp_dim=4, num_feat_per_dim=12, bandwidth=1.0)
    rf.freq, rf.shift, rf.scale = torch.from_numpy(g["pf_rf_freq"]), torch.from_numpy(g["pf_rf_shift"]), torch.tensor(float(g["pf_rf_scale"]))
    bounds = (np.array([-2.0, -1.0, -3.0, -2.0]), np.array([2.0, 1.0, 3.0, 2.0]))
    rbf_s = RBFFeat(num_feat_per_dim=5, bounds=bounds)
    rbf_j = RBFFeat(num_feat_per_dim=4, bounds=bounds, scale=0.7, state_wise_norm=False)
    np.testing.assert_allclose(rbf_s.centers.numpy(), g["pf_rbf_state_centers"], rtol=1e-6)
    np.testing.assert_allclose(rbf_s.scale.numpy(), g["pf_rbf_state_scale"], rtol=1e-6)
    stacks = dict(elem=[const_feat, functools.partial(sig_feat, scale=1.7), functools.partial(bell_feat, scale=0.6), sig_feat,
                        MultFeat((0, 2)), MultFeat((1, 2, 3)), ATan2Feat(0, 1), identity_feat],
                  rf=[rf], rbf_state=[rbf_s], rbf_joint=[rbf_j])
    for tag, fl in stacks.items():
        fs = FeatureStack(*fl)
        ref = g[f"pf_{tag}_feats"]
        assert fs.get_num_feat(4) == ref.shape[1]
        if tag == "elem":  # batched call of a mixed stack (the reference's MultFeat / ATan2Feat take 1-dim inputs only)
            np.testing.assert_allclose(fs(x).numpy(), ref, rtol=1e-5, atol=1e-6)
        one = torch.stack([torch.cat([f(x[i]).reshape(-1) for f in fl]) for i in range(x.shape[0])])
        np.testing.assert_allclose(one.numpy(), ref, rtol=1e-5, atol=1e-6)
        codes = fs.engine_codes(4)
        assert codes is not None and len(codes) == len(fl)
    assert FeatureStack(lambda t: t).engine_codes(4) is None and FeatureStack(RFFeat(3, 4, 1.0)).engine_codes(4) is None
    with pytest.raises(pyrado.TypeErr):
        MultFeat([0, 1])
    with pytest.raises(pyrado.ShapeErr):
        MultFeat((0,))
    with pytest.raises(pyrado.ValueErr):
        RBFFeat(1, bounds)
