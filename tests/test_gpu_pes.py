"""
GPU tests of parameter-exploration sampling (pyrado/sampling/parameter_exploration_sampler.py):
the grouped-weights launch (`sre_policy.group_size`: lane n evaluates parameter set n // group_size) against the oracle
and against one launch per parameter set, and the reference-facing ParameterExplorationSampler / ParameterSamplingResult
(Pyrado/tests/test_sampling.py:205-260 style: result structure, common domains / init states across the population,
determinism, agreement with sequential rollout()).
"""
import numpy as np
import pytest
import torch

from simurlacra_b200 import _lib as L
from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer
from simurlacra_b200.environment_wrappers import ActNormWrapper, DomainRandWrapperLive
from simurlacra_b200.environments.pysim.one_mass_oscillator import OneMassOscillatorSim
from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
from simurlacra_b200.policies import FNNPolicy, LinearPolicy
from simurlacra_b200.policies.features import FeatureStack, identity_feat
from simurlacra_b200.sampling import ParameterExplorationSampler, ParameterSamplingResult, rollout
from test_gpu_engine import DEV, _t

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


def _sets(P, sizes, seed=5, scale=0.4):
    nw = sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1))
    return (np.random.RandomState(seed).randn(P, nw) * scale).astype(np.float32)


@pytest.mark.parametrize("name,dt", [("qq-su", 1 / 250.0), ("qcp-su", 1 / 250.0)])
def test_grouped_fnn_matches_oracle(E, orc, name, dt):
    """fp64 build, 7 parameter sets x 9 lanes (+ a ragged last group), FNN[16,16]: every lane follows the oracle run
    with the same [n_sets, n_weights] matrix and group size."""
    kind = E.KIND_BY_NAME[name]
    d = E.dims(kind)
    R, P, H = 9, 7, 80
    N = P * R - 4
    sizes = [d["do"], 16, 16, d["da"]]
    w = _sets(P, sizes)
    cfg = E.make_cfg(kind, dt, H, dtype=torch.float64, act_norm=True)
    pol = E.make_policy(L.POL_FNN, _t(w, torch.float32), sizes=sizes, group_size=R)
    res = E.rollout(cfg, pol, N, H + 1, device=DEV, seed=2)
    ref = orc.rollout_batch(kind, dt, H, N, H, policy=orc.make_policy(orc.POL_FNN, w, sizes=sizes, group_size=R),
                            act_norm=True, seed=2)
    ln = res["length"].cpu().numpy()
    np.testing.assert_array_equal(ln, ref["length"])
    st, ac = res["state"].cpu().numpy(), res["act"].cpu().numpy()
    for n in range(N):
        T = ln[n]
        sc = np.maximum(1.0, np.abs(ref["state"][: T + 1, n]))
        assert (np.abs(st[: T + 1, n] - ref["state"][: T + 1, n]) / sc).max() < 1e-4
    np.testing.assert_allclose(ac[0], ref["act"][0], rtol=0, atol=2e-6)
    # the sets really differ: first actions of lanes with the same init state would otherwise coincide
    assert len(np.unique(np.round(ac[0, ::R, 0], 6))) == len(ac[0, ::R, 0])


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_grouped_launch_equals_one_launch_per_set(E, dtype):
    """Lanes [p R, (p+1) R) of the grouped launch == a launch of R lanes at env_offset p R with set p as the policy's
    only weights: same Philox streams, same policy arithmetic (bit-exact for the linear policy, which runs the same
    code on both routes; the FNN goes through the tiled kernel on the single-set route, so to round-off)."""
    R, P, H = 16, 6, 60
    cfg = E.make_cfg(L.QQ, 1 / 250.0, H, dtype=dtype, act_norm=True)
    from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer_qq

    distr = create_default_randomizer_qq().table(QQubeSwingUpSim._param_names)
    feats = ["const", "identity", "sin"]
    wl = (np.random.RandomState(1).randn(P, 1 + 2 * 6) * 0.1).astype(np.float32)
    sizes = [6, 16, 16, 1]
    wf = _sets(P, sizes)
    for kind, w, kw, exact in ((L.POL_LINEAR, wl, dict(feats=feats), True), (L.POL_FNN, wf, dict(sizes=sizes), False)):
        g = E.rollout(cfg, E.make_policy(kind, _t(w, torch.float32), group_size=R, **kw), P * R, H + 1, device=DEV,
                      distr=distr, seed=8)
        for p in range(P):
            s = E.rollout(cfg, E.make_policy(kind, _t(w[p], torch.float32), **kw), R, H + 1, device=DEV, distr=distr,
                          seed=8, env_offset=p * R)
            sl = slice(p * R, (p + 1) * R)
            assert torch.equal(g["params"][sl], s["params"])
            if exact:
                assert torch.equal(g["state"][:, sl], s["state"]) and torch.equal(g["ret"][sl], s["ret"])
            else:
                assert torch.equal(g["length"][sl], s["length"])
                # float32 policy on both builds (generic FNN / tanhf against the register-resident FNN16 / ex2-based tanh):
                # round-off amplified over 60 swing-up steps, measured against the largest state entry (velocities ~ 20)
                scale = max(1.0, float(s["state"].abs().max()))
                assert float((g["state"][:, sl] - s["state"]).abs().max()) < 1e-3 * scale


@pytest.mark.parametrize("sizes,prec,R", [([6, 64, 64, 1], L.PREC_TF32X3, 128), ([6, 64, 64, 1], L.PREC_TF32X3, 384),
                                          ([6, 64, 64, 1], L.PREC_FP32, 256), ([6, 16, 16, 1], L.PREC_FP32, 512),
                                          ([6, 24, 1], L.PREC_FP32, 128), ([6, 64, 64, 1], L.PREC_TF32X3, 96)])
def test_cta_aligned_groups_use_the_fast_variants(E, sizes, prec, R):
    """Groups that are multiples of the CTA size stage one parameter set per CTA and run the tensor-core / tiled /
    generic variants unchanged: bit-identical to one launch per set with that variant (R = 96 is not aligned and takes
    the per-lane route: equal to round-off).  The last group is ragged."""
    P, H = 5, 50
    N = P * R - 40
    cfg = E.make_cfg(L.QQ, 1 / 250.0, H, dtype=torch.float32, act_norm=True)
    from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer_qq

    distr = create_default_randomizer_qq().table(QQubeSwingUpSim._param_names)
    w = _sets(P, sizes, scale=0.2)
    g = E.rollout(cfg, E.make_policy(L.POL_FNN, _t(w, torch.float32), sizes=sizes, precision=prec, group_size=R), N, H + 1,
                  device=DEV, distr=distr, seed=4)
    assert int(g["status"].max()) == 0
    for p in range(P):
        n = min(R, N - p * R)
        s = E.rollout(cfg, E.make_policy(L.POL_FNN, _t(w[p], torch.float32), sizes=sizes, precision=prec), n, H + 1,
                      device=DEV, distr=distr, seed=4, env_offset=p * R)
        sl = slice(p * R, p * R + n)
        if R % 128 == 0 and (R % 256 == 0 or prec != L.PREC_FP32 or len(sizes) == 3):
            assert torch.equal(g["state"][:, sl], s["state"]) and torch.equal(g["ret"][sl], s["ret"])
        else:
            assert torch.equal(g["length"][sl], s["length"])
            assert float((g["state"][:, sl] - s["state"]).abs().max()) < 1e-3


def test_group_argument_errors(E):
    cfg = E.make_cfg(L.QQ, 1 / 250.0, 10, dtype=torch.float32)
    sizes = [6, 8, 1]
    w = _t(_sets(3, sizes), torch.float32)
    with pytest.raises(ValueError):
        E.make_policy(L.POL_FNN, w[0], sizes=sizes, group_size=4)  # grouped policies take a matrix
    with pytest.raises(L.EngineError):  # 4 groups of 4 lanes, 3 sets
        E.rollout(cfg, E.make_policy(L.POL_FNN, w, sizes=sizes, group_size=4), 16, 11, device=DEV)
    with pytest.raises(L.EngineError):
        p = E.make_policy(L.POL_FNN, w, sizes=sizes, group_size=4)
        p.group_size = -1
        E.rollout(cfg, p, 8, 11, device=DEV)
    # a [do,64,64,da] net asks for the tensor-core build; grouped weights take the per-lane route instead
    big = _t(_sets(2, [6, 64, 64, 1], scale=0.1), torch.float32)
    gb = E.rollout(cfg, E.make_policy(L.POL_FNN, big, sizes=[6, 64, 64, 1], group_size=4, precision=L.PREC_TF32X3), 8, 11,
                   device=DEV, seed=1)
    one = E.rollout(cfg, E.make_policy(L.POL_FNN, big[1], sizes=[6, 64, 64, 1]), 4, 11, device=DEV, seed=1, env_offset=4)
    assert float((gb["state"][:, 4:] - one["state"]).abs().max()) < 1e-4
    ok = E.rollout(cfg, E.make_policy(L.POL_FNN, w, sizes=sizes, group_size=4), 12, 11, device=DEV)
    assert int(ok["status"].max()) == 0


def _chain(max_steps=60):
    env = QQubeSwingUpSim(dt=1 / 250.0, max_steps=max_steps)
    return DomainRandWrapperLive(ActNormWrapper(env), create_default_randomizer(env))


def test_sampler_result_structure_and_common_random_numbers():
    env = _chain()
    torch.manual_seed(0)
    pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
    P, D, K = 5, 3, 2
    sets = pol.param_values.detach().unsqueeze(0) + 0.3 * torch.randn(P, pol.num_param)
    before = pol.param_values.detach().clone()
    pes = ParameterExplorationSampler(env, pol, num_init_states_per_domain=K, num_domains=D, num_workers=2, seed=7)
    assert pes.num_rollouts_per_param == D * K
    res = pes.sample(sets)
    assert isinstance(res, ParameterSamplingResult) and len(res) == P and res.num_rollouts == P * D * K
    assert torch.equal(res.parameters, sets) and res.mean_returns.shape == (P,)
    assert len(res.rollouts) == P and all(len(r) == D * K for r in res.rollouts)
    assert torch.equal(pol.param_values.detach(), before)  # the module keeps its own parameters
    for p, s in enumerate(res):
        assert s.num_rollouts == D * K and torch.equal(s.params, sets[p])
        np.testing.assert_allclose(s.mean_undiscounted_return, res.mean_returns[p], rtol=1e-5)
        for r, ro in enumerate(s.rollouts):
            # every parameter set sees the same domains and initial states
            assert ro.rollout_info["domain_param"] == res[0].rollouts[r].rollout_info["domain_param"]
            np.testing.assert_array_equal(ro.states[0], res[0].rollouts[r].states[0])
    # D distinct domains, each used for K rollouts (the reference's `domain_params *= K` ordering)
    masses = [ro.rollout_info["domain_param"]["mass_pend_pole"] for ro in res[0].rollouts]
    assert len(set(masses)) == D and masses[:D] == masses[D:]
    assert len({tuple(ro.states[0]) for ro in res[0].rollouts}) == D * K
    sub = res[1:3]
    assert isinstance(sub, ParameterSamplingResult) and len(sub) == 2
    np.testing.assert_array_equal(sub.mean_returns, res.mean_returns[1:3])
    # determinism: same seed and call count -> same numbers; the next call draws new domains
    again = ParameterExplorationSampler(env, pol, num_init_states_per_domain=K, num_domains=D, seed=7).sample(sets)
    np.testing.assert_array_equal(again.mean_returns, res.mean_returns)
    nxt = pes.sample(sets)
    assert nxt[0].rollouts[0].rollout_info["domain_param"] != res[0].rollouts[0].rollout_info["domain_param"]
    # returns only
    lean = ParameterExplorationSampler(env, pol, num_init_states_per_domain=K, num_domains=D, seed=7).sample(
        sets, return_rollouts=False)
    np.testing.assert_array_equal(lean.mean_returns, res.mean_returns)
    assert lean[0].rollouts == []


def test_sampler_agrees_with_sequential_rollouts():
    """Every rollout of the result == rollout() of the policy carrying that parameter set on the env without the DR
    wrapper, from the recorded domain parameters and initial state."""
    env = _chain(max_steps=40)
    torch.manual_seed(3)
    pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
    sets = 0.5 * torch.randn(3, pol.num_param)
    pes = ParameterExplorationSampler(env, pol, num_init_states_per_domain=1, num_domains=2, seed=1)
    res = pes.sample(sets)
    for s in res:
        pol.param_values = s.params
        for ro in s.rollouts:
            ref = rollout(pes.env, pol, reset_kwargs=dict(init_state=ro.states[0], domain_param=ro.rollout_info["domain_param"]))
            assert len(ref) == len(ro)
            np.testing.assert_allclose(ref.states, ro.states, rtol=0, atol=1e-4)
            np.testing.assert_allclose(ref.undiscounted_return(), ro.undiscounted_return(), rtol=1e-4)
    # fixed initial states, no DR wrapper, linear policy
    env2 = OneMassOscillatorSim(dt=0.01, max_steps=50)
    lin = LinearPolicy(env2.spec, FeatureStack(identity_feat))
    inits = [np.array([0.3, 0.0]), np.array([-0.2, 0.1])]
    pes2 = ParameterExplorationSampler(env2, lin, num_init_states_per_domain=2, num_domains=1, seed=0)
    r2 = pes2.sample(torch.tensor([[-1.0, -0.5], [0.5, 0.0], [0.0, 0.0]]), init_states=inits)
    assert len(r2) == 3 and r2.num_rollouts == 6
    for s in r2:
        for ro, s0 in zip(s.rollouts, inits):
            np.testing.assert_allclose(ro.states[0], s0, atol=1e-7)
    assert not np.allclose(r2[0].rollouts[0].states[-1], r2[1].rollouts[0].states[-1])
    with pytest.raises(RuntimeError):
        pes2.sample(torch.zeros(2, 2), init_states=inits[:1])
    with pytest.raises(RuntimeError):
        pes2.sample(torch.zeros(2, 5))


def test_sampler_generic_path():
    """A policy the kernels do not implement (softplus hidden layer) takes the reference's loop, same result type."""
    env = _chain(max_steps=15)
    pol = FNNPolicy(env.spec, hidden_sizes=[8], hidden_nonlin=torch.nn.functional.softplus)
    assert pol.engine_desc(torch.device("cuda", 0)) is None
    sets = 0.2 * torch.randn(2, pol.num_param)
    res = ParameterExplorationSampler(env, pol, num_init_states_per_domain=1, num_domains=2, seed=0).sample(sets)
    assert len(res) == 2 and res.num_rollouts == 4 and res.mean_returns.shape == (2,)
    assert res[0].rollouts[0].rollout_info["domain_param"] == res[1].rollouts[0].rollout_info["domain_param"]


def test_population_shards_give_the_same_rollouts():
    """Engine extension (one process per GPU): rank r of w rolls out its block of the parameter sets; mean returns and
    rollouts of any split equal the unsharded call (lanes keyed on the global index, domains / init states on the seed)."""
    env = _chain(max_steps=40)
    torch.manual_seed(5)
    pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
    P = 7
    sets = 0.4 * torch.randn(P, pol.num_param)
    mk = lambda shard: ParameterExplorationSampler(env, pol, num_init_states_per_domain=2, num_domains=2, seed=11, shard=shard)
    full = mk(None).sample(sets)
    for ws in (2, 3):
        merged = np.full(P, np.nan)
        for rank in range(ws):
            part = mk((rank, ws)).sample(sets)
            own = ~np.isnan(part.mean_returns)
            assert own.sum() in (P // ws, P // ws + 1) and np.isnan(merged[own]).all()
            merged[own] = part.mean_returns[own]
            for p in np.flatnonzero(own):
                assert len(part[p].rollouts) == 4
                for a, b in zip(part[p].rollouts, full[p].rollouts):
                    np.testing.assert_array_equal(a.states, b.states)
                    assert a.rollout_info["domain_param"] == b.rollout_info["domain_param"]
            assert all(part[p].rollouts == [] for p in np.flatnonzero(~own))
        np.testing.assert_array_equal(merged, full.mean_returns)


_PES_WORKER = """
import os, sys
sys.path.insert(0, os.environ["SRE_ROOT"])
import numpy as np, torch, torch.distributed as dist
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer
from simurlacra_b200.environment_wrappers import ActNormWrapper, DomainRandWrapperLive
from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
from simurlacra_b200.policies import FNNPolicy
from simurlacra_b200.sampling import ParameterExplorationSampler
sim = QQubeSwingUpSim(dt=1 / 250.0, max_steps=40)
env = DomainRandWrapperLive(ActNormWrapper(sim), create_default_randomizer(sim))
torch.manual_seed(5)
pol = FNNPolicy(env.spec, hidden_sizes=[16, 16], hidden_nonlin=torch.tanh)
sets = 0.4 * torch.randn(9, pol.num_param)
both = ParameterExplorationSampler(env, pol, 2, 2, seed=3).sample(sets, return_rollouts=False)        # sharded + all-gather
alone = ParameterExplorationSampler(env, pol, 2, 2, seed=3, shard=(0, 1)).sample(sets, return_rollouts=False)
assert not np.isnan(both.mean_returns).any()
np.testing.assert_array_equal(both.mean_returns, alone.mean_returns)
dist.barrier()
sys.stdout.write(f"rank {dist.get_rank()} ok\\n")
dist.destroy_process_group()
"""


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_population_sharding_over_nccl(tmp_path):
    import os, socket, subprocess, sys

    script = tmp_path / "worker.py"
    script.write_text(_PES_WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       capture_output=True, text=True, env=dict(os.environ, SRE_ROOT=root), timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout
