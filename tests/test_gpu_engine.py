"""
GPU tests of the fused rollout path against the CPU oracle on the same seeded inputs: Philox domain
randomization, init-state sampling, observation noise, fused FNN / Linear policies, ActNorm, trajectory records,
auto-reset sampling, statistics, shard invariance.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


def _t(a, dtype):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(DEV).contiguous()


def default_distr(name):
    """The reference's default randomizer of env `name` as table entries (tests/golden/dr_reference.npz)."""
    g = golden("dr_reference")
    keys = [str(k) for k in g[f"{name}_keys"]]
    fn = {"qq-su": "traj_qq-su", "qcp-su": "traj_qcp-su", "qbb": "traj_qbb", "bob": "traj_bob", "omo": "traj_omo"}[name]
    names = [str(k) for k in golden(fn)["param_keys"]]
    ent = []
    for k, r in zip(keys, g[f"{name}_table"]):
        ent.append(dict(param_index=names.index(k), kind="normal" if r[0] == 0 else "uniform", a=r[1], b=r[2],
                        clip_lo=r[4], clip_up=r[5]))
    return ent


ENVS = [("qq-su", 1 / 250.0, 600, {}), ("qcp-su", 1 / 250.0, 600, dict(long_pole=True)), ("qbb", 0.01, 500, {}),
        ("bob", 0.01, 500, {}), ("omo", 0.02, 300, {})]


@pytest.mark.parametrize("name,dt,H,kw", ENVS)
def test_dr_and_init_sampling_match_oracle(E, orc, name, dt, H, kw):
    """sre_sample_params / sre_sample_init against the oracle's Philox restatement.  The normal draw is evaluated in
    float32 on the device from exactly representable inputs and in fp64 by the oracle (same expression): the float32
    parameter agrees to 2 ulp (most draws exactly); uniform / bernoulli draws are bit-exact; init states to 1e-15."""
    kind = E.KIND_BY_NAME[name]
    d = E.dims(kind)
    ent = default_distr(name)
    cfg = E.make_cfg(kind, dt, H, dtype=torch.float64, **kw)
    N = 512
    params = _t(np.tile(E.nominal_params(kind, kw.get("long_pole", False)), (N, 1)), torch.float64)
    E.sample_params(cfg, E.make_distr_table(ent), params, seed=11, env_offset=1000, episode=3)
    tab = orc.make_distr_table(ent)
    ref = np.tile(orc.nominal_params(kind, kw.get("long_pole", False)), (N, 1))
    for n in range(N):
        for i, e in enumerate(ent):
            ref[n, e["param_index"]] = orc.lib().orc_sample_entry(C.byref(tab[i]), 11, 1000 + n, 3, i)
    got = params.cpu().numpy()
    # normal draws: |z_f32 - z_f64| <= 3 ulp(z), so the parameter a + b z agrees to 1e-6 of its scale (|a| + |b|)
    err = np.abs(np.float32(got) - np.float32(ref))
    for e in ent:
        j = e["param_index"]
        if e["kind"] == "normal":
            assert (err[:, j] <= 1e-6 * (abs(e["a"]) + abs(e["b"]))).all(), (j, err[:, j].max())
            assert (err[:, j] == 0).mean() > 0.5
        else:
            assert (err[:, j] == 0).all()
    # init states (default init space, parameters from above)
    init = E.sample_init(cfg, params, N, seed=11, env_offset=1000, episode=3).cpu().numpy()
    e = orc.OracleEnv(kind, dt, H, long_pole=kw.get("long_pole", False))
    for n in range(0, N, 7):
        e.set_params(got[n])
        r = e.sample_init(11, 1000 + n, 3)
        np.testing.assert_allclose(init[n], r, rtol=1e-13, atol=1e-15)


@pytest.mark.parametrize("name,dt,H,kw", ENVS)
@pytest.mark.parametrize("policy", ["idle", "fnn16", "linear"])
def test_rollout_fp64_matches_oracle(E, orc, name, dt, H, kw, policy):
    """Whole fused path in fp64 (DR + sampled init + policy + ActNorm + obs noise + records) against
    orc_rollout_batch.  Parameters are injected from the engine's own DR draw so that 1-ulp float32 sampling
    differences cannot leak into the comparison."""
    from simurlacra_b200 import _lib as L

    kind = E.KIND_BY_NAME[name]
    d = E.dims(kind)
    N, Hs = 96, min(H, 120)
    rng = np.random.RandomState(5)
    lp = kw.get("long_pole", False)
    cfg = E.make_cfg(kind, dt, Hs, dtype=torch.float64, act_norm=True, noise_std=0.01 * np.ones(d["do"]),
                     noise_mean=0.001 * np.ones(d["do"]), **kw)
    params = _t(np.tile(E.nominal_params(kind, lp), (N, 1)), torch.float64)
    E.sample_params(cfg, E.make_distr_table(default_distr(name)), params, seed=7)
    if policy == "idle":
        pol, opol = E.make_policy(L.POL_IDLE), orc.make_policy(orc.POL_IDLE)
    elif policy == "fnn16":
        sizes = [d["do"], 16, 16, d["da"]]
        w = (rng.randn(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(3))) * 0.4).astype(np.float32)
        pol, opol = E.make_policy(L.POL_FNN, _t(w, torch.float32), sizes=sizes), orc.make_policy(orc.POL_FNN, w, sizes=sizes)
    else:
        feats = ["const", "identity", "squared", "sin"]
        # small weights: squared velocity features reach 1e3, and float32 cancellation in the policy is amplified by
        # the closed loop
        w = (rng.randn(d["da"] * (1 + 3 * d["do"])) * 0.02).astype(np.float32)
        pol, opol = E.make_policy(L.POL_LINEAR, _t(w, torch.float32), feats=feats), orc.make_policy(orc.POL_LINEAR, w, feats=feats)
    res = E.rollout(cfg, pol, N, Hs + 1, device=DEV, params=params, seed=7, env_offset=40)
    ref = orc.rollout_batch(kind, dt, Hs, N, Hs, policy=opol, long_pole=lp, act_norm=True,
                            noise_std=0.01 * np.ones(d["do"]), noise_mean=0.001 * np.ones(d["do"]),
                            params=params.cpu().numpy(), seed=7, env_offset=40)
    ln = res["length"].cpu().numpy()
    np.testing.assert_array_equal(ln, ref["length"])  # lengths bit-exact
    st, ob, ac, rw = (res[k].cpu().numpy() for k in ("state", "obs", "act", "rew"))
    # with a float32 policy in the loop (FMA on the device, separate roundings + libm tanhf in the oracle) the
    # closed loop agrees to the fp32 tolerance of north_star (1e-4); without a policy to fp64 round-off
    tol = 1e-9 if policy == "idle" else 1e-4
    for n in range(N):
        T = ln[n]
        sc = np.maximum(1.0, np.abs(ref["state"][: T + 1, n]))
        assert (np.abs(st[: T + 1, n] - ref["state"][: T + 1, n]) / sc).max() < tol
        so = np.maximum(1.0, np.abs(ref["obs"][: T + 1, n]))
        assert (np.abs(ob[: T + 1, n] - ref["obs"][: T + 1, n]) / so).max() < tol
        assert np.abs(ac[:T, n] - ref["act"][:T, n]).max() < max(tol, 1e-5) * max(1.0, np.abs(ref["act"][:T, n]).max())
        np.testing.assert_allclose(rw[:T, n], ref["rew"][:T, n], rtol=10 * tol, atol=10 * tol)
    np.testing.assert_allclose(res["ret"].cpu().numpy(), ref["ret"], rtol=100 * tol, atol=100 * tol)
    stats = res["stats"].cpu().numpy()
    assert stats[L.STAT_STEPS] == ln.sum()
    assert stats[L.STAT_EPISODES] == (ref["done"].sum(0) > 0).sum()


def test_policy_forward_golden(E):
    from simurlacra_b200 import _lib as L

    g = golden("policy_forward")
    for tag in "abcd":
        sizes = [int(s) for s in g[f"fnn_{tag}_sizes"]]
        pol = E.make_policy(L.POL_FNN, _t(g[f"fnn_{tag}_weights"], torch.float32), sizes=sizes)
        y = E.policy_forward(pol, _t(g[f"fnn_{tag}_x"], torch.float32), sizes[-1]).cpu().numpy()
        np.testing.assert_allclose(y, g[f"fnn_{tag}_y"], rtol=2e-5, atol=2e-6)
    pol = E.make_policy(L.POL_LINEAR, _t(g["lin_W"].reshape(-1), torch.float32), feats=[str(f) for f in g["lin_feats"]])
    y = E.policy_forward(pol, _t(g["lin_x"], torch.float32), 2).cpu().numpy()
    np.testing.assert_allclose(y, g["lin_y"], rtol=2e-5, atol=2e-5)


def test_parametrised_linear_features_golden(E):
    """sig / bell / MultFeat / ATan2Feat / RFFeat / RBFFeat inside the Linear-policy kernel (csrc/sre_policy.cuh
    linear_accumulate) against W @ feats of the unmodified reference (policies/features.py:157-161, :181-445), through the
    LinearPolicy class and as a fused policy inside a rollout launch."""
    import functools

    from simurlacra_b200 import _lib as L
    from simurlacra_b200.policies.features import (ATan2Feat, FeatureStack, MultFeat, RBFFeat, RFFeat, bell_feat, const_feat,
                                                   identity_feat, sig_feat)

    g = golden("policy_forward")
    rf = RFFeat(inp_dim=4, num_feat_per_dim=12, bandwidth=1.0)
    rf.freq, rf.shift, rf.scale = torch.from_numpy(g["pf_rf_freq"]), torch.from_numpy(g["pf_rf_shift"]), torch.tensor(float(g["pf_rf_scale"]))
    bounds = (np.array([-2.0, -1.0, -3.0, -2.0]), np.array([2.0, 1.0, 3.0, 2.0]))
    stacks = dict(elem=[const_feat, functools.partial(sig_feat, scale=1.7), functools.partial(bell_feat, scale=0.6), sig_feat,
                        MultFeat((0, 2)), MultFeat((1, 2, 3)), ATan2Feat(0, 1), identity_feat],
                  rf=[rf], rbf_state=[RBFFeat(num_feat_per_dim=5, bounds=bounds)],
                  rbf_joint=[RBFFeat(num_feat_per_dim=4, bounds=bounds, scale=0.7, state_wise_norm=False)])
    x = _t(g["pf_x"], torch.float32)
    for tag, fl in stacks.items():
        codes = FeatureStack(*fl).engine_codes(4)
        pol = E.make_policy(L.POL_LINEAR, _t(g[f"pf_{tag}_W"].reshape(-1), torch.float32), feats=codes)
        y = E.policy_forward(pol, x, 2).cpu().numpy()
        np.testing.assert_allclose(y, g[f"pf_{tag}_y"], rtol=3e-5, atol=3e-5, err_msg=tag)
    # inside a rollout launch: OneMassOscillator (2 observations, 1 action) with an RBF + ATan2 stack, fused against stepping
    from simurlacra_b200.environments.pysim.one_mass_oscillator import OneMassOscillatorSim
    from simurlacra_b200.policies.feed_back.linear import LinearPolicy
    from simurlacra_b200.sampling import rollout
    from simurlacra_b200.sampling.rollout import compile_chain

    env = OneMassOscillatorSim(dt=0.02, max_steps=40)
    torch.manual_seed(2)
    pol = LinearPolicy(env.spec, FeatureStack(const_feat, RBFFeat(3, bounds=(np.array([-1.0, -2.0]), np.array([1.0, 2.0]))),
                                              ATan2Feat(0, 1), functools.partial(bell_feat, scale=0.5)))
    assert pol.engine_desc(env.device) is not None and compile_chain(env) is not None
    fused = rollout(env, pol, seed=1, reset_kwargs=dict(init_state=np.array([0.3, -0.2])))
    loop = rollout(env, pol, seed=1, record_dts=True, reset_kwargs=dict(init_state=np.array([0.3, -0.2])))
    np.testing.assert_allclose(fused.actions, loop.actions, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(fused.observations, loop.observations, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_fnn_closed_loop_golden(E, dtype):
    """Config C1: QQube + FNN in closed loop against the reference run (torch CPU policy, numpy env); covers the
    register-resident fast paths [64,64], [32,32], [16,16] and ActNorm. 1e-4 over the horizon."""
    from simurlacra_b200 import _lib as L

    g = golden("fnn_closed_loop_qq")
    for tag in ("h64", "h16_actnorm", "h32"):
        sizes = [int(s) for s in g[f"{tag}_sizes"]]
        cfg = E.make_cfg(L.QQ, 1 / 250.0, 600, dtype=dtype, act_norm=bool(g[f"{tag}_actnorm"]))
        pol = E.make_policy(L.POL_FNN, _t(g[f"{tag}_weights"], torch.float32), sizes=sizes)
        res = E.rollout(cfg, pol, 1, 601, device=DEV, init=_t(g[f"{tag}_init"][None, :], dtype))
        S = g[f"{tag}_states"]
        H = S.shape[0] - 1
        assert int(res["length"][0]) == H
        err = np.abs(res["state"][: H + 1, 0].double().cpu().numpy() - S) / np.maximum(1.0, np.abs(S))
        assert err.max() < (1e-4 if dtype == torch.float64 else 2e-3), (tag, err.max())
        assert np.abs(res["state"][:101, 0].double().cpu().numpy() - S[:101]).max() < 1e-4


def test_qbb_kin_golden(E):
    """reset() of the ball balancer runs the 100-iteration float32 SGD inverse kinematics on the device."""
    from simurlacra_b200 import _lib as L

    rows = golden("qbb_kin")["rows"]
    N = len(rows)
    cfg = E.make_cfg(L.QBB, 0.01, 500, dtype=torch.float64)
    params = np.tile(E.nominal_params(L.QBB), (N, 1))
    params[:, 4], params[:, 3] = rows[:, 0], rows[:, 1]
    init = np.zeros((N, 8))
    init[:, 0] = rows[:, 2]  # th_x; offsets are zero in the nominal parameters
    state = torch.empty(N, 8, dtype=torch.float64, device=DEV)
    hidden = torch.empty(N, 2, dtype=torch.float64, device=DEV)
    curr = torch.empty(N, dtype=torch.int32, device=DEV)
    E.reset(cfg, _t(params, torch.float64), _t(init, torch.float64), state, hidden, curr)
    assert np.abs(hidden[:, 0].cpu().numpy() - rows[:, 3]).max() < 2e-6
