"""
Pins the CPU oracle (oracle/sre_oracle.c) against
  * the trajectories produced by the unmodified reference (tests/golden/*.npz, made by oracle/gen_golden.py),
  * the reference's own known-answer tests (Pyrado/tests/test_tasks.py:82-99),
  * the Random123 known-answer vectors for Philox4x32-10.
All CPU, runs in seconds.
"""
import numpy as np
import pytest

from conftest import ENV_FILES, FNN64_CASES, OMO_TASK, STAB_TASKS, golden


def _make(O, kind_name, dt, max_steps, kw):
    kind = O.KIND_BY_NAME[kind_name]
    kw = dict(kw)
    stab = kw.pop("stab", None)
    e = O.OracleEnv(kind, dt, max_steps, stab=(stab == "qcp-st"), **kw)
    if stab:
        t = STAB_TASKS[stab]
        e.set_task(t["state_des"], t["Q"], t["R"], O.REW_EXP if t["rew"] == "exp" else O.REW_QUADR, radial_mask=0b10,
                   final_flags=(O.FR_ENABLED | O.FR_STATE_DEP | O.FR_TIME_DEP) if t["final"] else 0, final_factor=1e3)
    if kind_name == "omo":  # the golden OMO env was built with task_args=dict(state_des=[0.5, 0])
        e.set_task(OMO_TASK["state_des"], OMO_TASK["Q"], OMO_TASK["R"], O.REW_QUADR,
                   final_flags=O.FR_ENABLED | O.FR_ALWAYS_NEG, final_factor=1e3)
    return e


@pytest.mark.parametrize("fname,kind_name,kw", ENV_FILES)
def test_trajectories_match_reference(orc, fname, kind_name, kw):
    g = golden(fname)
    for c in range(int(g["n_cases"])):
        e = _make(orc, kind_name, float(g["dt"]), int(g["max_steps"]), kw)
        e.set_params(g[f"c{c}_params"])
        lo, up, alo, aup = e.spaces()
        np.testing.assert_allclose(lo, g[f"c{c}_state_lo"], rtol=1e-15)
        np.testing.assert_allclose(up, g[f"c{c}_state_up"], rtol=1e-15)
        np.testing.assert_allclose(alo, g[f"c{c}_act_lo"], rtol=1e-15)
        np.testing.assert_allclose(aup, g[f"c{c}_act_up"], rtol=1e-15)
        obs = e.reset(g[f"c{c}_init"])
        np.testing.assert_allclose(obs, g[f"c{c}_obs"][0], rtol=1e-14, atol=1e-15)
        if g[f"c{c}_hidden"].shape[1]:
            np.testing.assert_allclose(e.hidden, g[f"c{c}_hidden"][0], rtol=0, atol=1e-7)
        acts, S, R, D = g[f"c{c}_actions"], g[f"c{c}_states"], g[f"c{c}_rews"], g[f"c{c}_dones"]
        assert len(acts) == int(g[f"c{c}_length"])
        for t in range(len(acts)):
            obs, r, d = e.step(acts[t].copy())
            scale = np.maximum(1.0, np.abs(S[t + 1]))
            assert np.all(np.abs(e.state - S[t + 1]) <= 1e-11 * scale), (fname, c, t)
            assert abs(r - R[t]) <= 1e-11 * max(1.0, abs(R[t])), (fname, c, t)
            assert bool(d) == bool(D[t]), (fname, c, t)  # done flags bit-exact
            np.testing.assert_allclose(obs, g[f"c{c}_obs"][t + 1], rtol=1e-10, atol=1e-11)
        assert bool(D[-1])  # every golden episode ends (state bound or max_steps)


def test_modulated_rew_known_answer(orc):
    """Pyrado/tests/test_tasks.py:82-99: s=[1,2,3,4], Q=I, R=I, a=0."""
    e = orc.OracleEnv(orc.QQ, 0.01, 10)
    e.set_task(np.zeros(4), np.eye(4), np.eye(1), orc.REW_QUADR, radial_mask=0b1011)
    e.set_radial_mod([2.0, 2.0, 2.0, 2.0])
    assert e.step_rew([1.0, 2.0, 3.0, 4.0], [0.0]) == -(1 ** 2 + 3 ** 2)
    e.set_task(np.zeros(4), np.eye(4), np.eye(1), orc.REW_QUADR, radial_mask=0b1010)
    e.set_radial_mod([0.0, 2.0, 0.0, 3.0])
    assert e.step_rew([1.0, 2.0, 3.0, 4.0], [0.0]) == -(1 ** 2 + 3 ** 2 + 1 ** 2)


def test_philox_known_answers(orc):
    """Random123 kat_vectors, philox4x32 10 rounds."""
    kat = [
        ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
        ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
        ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
         [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
    ]
    for ctr, key, exp in kat:
        assert list(orc.philox(ctr, key)) == exp


def test_policy_forward_matches_torch(orc):
    g = golden("policy_forward")
    for tag in "abcd":
        sizes = [int(s) for s in g[f"fnn_{tag}_sizes"]]
        pol = orc.make_policy(orc.POL_FNN, g[f"fnn_{tag}_weights"], sizes=sizes)
        x, y = g[f"fnn_{tag}_x"], g[f"fnn_{tag}_y"]
        for i in range(x.shape[0]):
            out = orc.policy_forward(pol, x[i], sizes[-1])
            np.testing.assert_allclose(out, y[i], rtol=2e-5, atol=2e-6)
    feats = [str(f) for f in g["lin_feats"]]
    pol = orc.make_policy(orc.POL_LINEAR, g["lin_W"], feats=feats)
    for i in range(g["lin_x"].shape[0]):
        out = orc.policy_forward(pol, g["lin_x"][i], 2)
        np.testing.assert_allclose(out, g["lin_y"][i], rtol=2e-5, atol=2e-5)


def test_qbb_kin_matches_reference(orc):
    rows = golden("qbb_kin")["rows"]
    for r_arm, l_plate, th, ang in rows:
        assert abs(orc.qbb_kin(r_arm, l_plate, th) - ang) < 2e-6


def test_fnn_closed_loop(orc):
    """Config C1: QQube + FNN in closed loop; the oracle's fp32 policy differs from torch's GEMM summation order
    by ~1e-7 per action, so the comparison is toleranced (1e-4 over the horizon, north_star)."""
    g = golden("fnn_closed_loop_qq")
    for tag in ("h64", "h16_actnorm", "h32"):
        sizes = [int(s) for s in g[f"{tag}_sizes"]]
        pol = orc.make_policy(orc.POL_FNN, g[f"{tag}_weights"], sizes=sizes)
        S = g[f"{tag}_states"]
        H = S.shape[0] - 1
        res = orc.rollout_batch(orc.QQ, 1 / 250.0, 600, 1, 600, policy=pol, act_norm=bool(g[f"{tag}_actnorm"]),
                                init=g[f"{tag}_init"][None, :])
        assert int(res["length"][0]) == H
        err = np.abs(res["state"][: H + 1, 0] - S) / np.maximum(1.0, np.abs(S))
        assert err.max() < 1e-4, (tag, err.max())
        np.testing.assert_allclose(res["rew"][:H, 0], g[f"{tag}_rews"], rtol=1e-4, atol=1e-6)
        np.testing.assert_allclose(res["act"][:H, 0], g[f"{tag}_acts"], rtol=1e-3, atol=1e-4)


@pytest.mark.parametrize("tag", list(FNN64_CASES))
def test_fnn64_closed_loop_envs(orc, tag):
    """FNN [do, 64, 64, da] in closed loop on QQube / QCartPole / QBallBalancer / Pendulum against the reference run
    (torch float32 policy, numpy float64 env): states 1e-4 over the whole episode, lengths equal."""
    g = golden("fnn64_closed_loop_envs")
    kind_name, kw = FNN64_CASES[tag]
    kind = orc.KIND_BY_NAME[kind_name]
    sizes = [int(s) for s in g[f"{tag}_sizes"]]
    pol = orc.make_policy(orc.POL_FNN, g[f"{tag}_weights"], sizes=sizes)
    S = g[f"{tag}_states"]
    H, ms = S.shape[0] - 1, int(g[f"{tag}_max_steps"])
    init = g[f"{tag}_init"][None, :]
    res = orc.rollout_batch(kind, float(g[f"{tag}_dt"]), ms, 1, ms, policy=pol, act_norm=bool(g[f"{tag}_actnorm"]),
                            init=init, **kw)
    assert int(res["length"][0]) == H
    err = np.abs(res["state"][: H + 1, 0] - S) / np.maximum(1.0, np.abs(S))
    assert err.max() < 1e-4, (tag, err.max())
    np.testing.assert_allclose(res["rew"][:H, 0], g[f"{tag}_rews"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(res["obs"][: H + 1, 0], g[f"{tag}_obs"], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("simple", [False, True])
def test_step_diff_state_values(orc, simple):
    g = golden("step_diff_qcp")
    key = f"{'simple' if simple else 'full'}_f64"
    s0, acts = g[f"{key}_s0"], g[f"{key}_acts"]
    e = orc.OracleEnv(orc.QCP, 1 / 20.0, 300, simple_dynamics=simple, long_pole=True)
    B, H = s0.shape[0], acts.shape[0]
    for b in range(B):
        s, thdd = s0[b].copy(), 0.0
        for t in range(H):
            # wrappers: ObsActCat -> ActNorm de-normalises act and act_diff=0 (6 * a for +-6 V)
            ns, r, d, thdd = e.step_diff_state(s, 6.0 * acts[t, b, 0], 0.0, thdd)
            np.testing.assert_allclose(ns, g[f"{key}_states"][t, b], rtol=1e-9, atol=1e-10)
            assert abs(r - g[f"{key}_rews"][t, b]) < 1e-10
            assert abs(thdd - g[f"{key}_thdd"][t, b]) < 1e-8 * max(1.0, abs(thdd))
            assert d == bool(g[f"{key}_dones"][t, b])
            s = ns


def test_dr_sampling_moments(orc):
    """The engine's Philox sampler cannot reproduce torch's Mersenne-Twister stream; it must reproduce the
    distributions the reference's default randomizers define (default_randomizers.py) incl. clipping."""
    g = golden("dr_reference")
    for name in ("qq-su", "qcp-su", "qbb", "bob", "omo"):
        kind = orc.KIND_BY_NAME[name]
        keys = [str(k) for k in g[f"{name}_keys"]]
        table = g[f"{name}_table"]
        ent = []
        for i, row in enumerate(table):
            ent.append(dict(param_index=i % orc.dims(kind)["dp"], kind="normal" if row[0] == 0 else "uniform",
                            a=row[1], b=row[2], clip_lo=row[4], clip_up=row[5]))
        arr = orc.make_distr_table(ent)
        n = 4000
        import ctypes as C
        draws = np.array([[orc.lib().orc_sample_entry(C.byref(arr[i]), 0, env, 0, i) for i in range(len(ent))]
                          for env in range(n)], dtype=np.float64)
        for i, row in enumerate(table):
            ref_mean, ref_std = g[f"{name}_mean"][i], g[f"{name}_std"][i]
            se = ref_std / np.sqrt(2000) + ref_std / np.sqrt(n)
            assert abs(draws[:, i].mean() - ref_mean) <= 5 * se + 1e-12, (name, keys[i])
            if ref_std > 0:
                assert abs(draws[:, i].std() / ref_std - 1) < 0.12, (name, keys[i])
            # samples are float32 like the reference's: the clip bounds apply after rounding to float32
            assert draws[:, i].min() >= np.float32(row[4]) and draws[:, i].max() <= np.float32(row[5])


def test_rollout_batch_deterministic_and_split_invariant(orc):
    """Analogue of Pyrado/tests/test_sampling.py:594-646: same seed => same rollouts for any partitioning."""
    g = golden("dr_reference")
    table = g["qq-su_table"]
    keys = [str(k) for k in g["qq-su_keys"]]
    names = [str(k) for k in golden("traj_qq-su")["param_keys"]]
    ent = [dict(param_index=names.index(k), kind="normal", a=r[1], b=r[2], clip_lo=r[4], clip_up=r[5])
           for k, r in zip(keys, table)]
    arr = orc.make_distr_table(ent)
    pol = orc.make_policy(orc.POL_IDLE)
    full = orc.rollout_batch(orc.QQ, 1 / 250.0, 50, 16, 50, policy=pol, distr=arr, seed=3, record=False)
    a = orc.rollout_batch(orc.QQ, 1 / 250.0, 50, 8, 50, policy=pol, distr=arr, seed=3, record=False)
    b = orc.rollout_batch(orc.QQ, 1 / 250.0, 50, 8, 50, policy=pol, distr=arr, seed=3, env_offset=8, record=False)
    np.testing.assert_array_equal(full["ret"], np.concatenate([a["ret"], b["ret"]]))
    np.testing.assert_array_equal(full["params"], np.concatenate([a["params"], b["params"]]))
    assert len(np.unique(full["ret"])) == 16  # rollouts within one batch differ


def test_grouped_policy_weights(orc):
    """Parameter-exploration layout (parameter_exploration_sampler.py:302-313: every parameter set runs the same
    rollouts): lanes [p R, (p+1) R) of a grouped batch == a batch of R lanes at env_offset p R with set p alone."""
    R, P, H = 3, 4, 40
    sizes = [6, 8, 1]
    rng = np.random.RandomState(0)
    w = (rng.randn(P, 6 * 8 + 8 + 8 + 1) * 0.4).astype(np.float32)
    g = orc.rollout_batch(orc.QQ, 1 / 250.0, H, P * R, H, policy=orc.make_policy(orc.POL_FNN, w, sizes=sizes, group_size=R),
                          act_norm=True, seed=5)
    for p in range(P):
        s = orc.rollout_batch(orc.QQ, 1 / 250.0, H, R, H, policy=orc.make_policy(orc.POL_FNN, w[p], sizes=sizes),
                              act_norm=True, seed=5, env_offset=p * R)
        np.testing.assert_array_equal(g["state"][:, p * R:(p + 1) * R], s["state"])
        np.testing.assert_array_equal(g["ret"][p * R:(p + 1) * R], s["ret"])
    assert len(np.unique(g["act"][0, ::R, 0])) == P


def test_returns_match_reference(orc):
    """orc_returns against discounted_value / gae_returns of the unmodified reference (tests/golden/returns_reference.npz,
    generated by oracle/gen_golden.py): single complete rollouts, then all of them chained in one lane (auto-reset
    layout) next to an idle lane -- every episode must come out unchanged."""
    g = golden("returns_reference")
    rews, flags, vals, rets, advs = [], [], [], [], []
    for c in range(int(g["n_cases"])):
        r, v = g[f"c{c}_rew"], g[f"c{c}_val"]
        gamma, lamb = g[f"c{c}_gamma_lamb"]
        n = len(r)
        flag = np.array([0] * (n - 1) + [1, 2], dtype=np.uint8)
        ret, adv = orc.returns(np.append(r, 0.0)[:, None], flag[:, None], gamma, values=v[:, None], lamb=lamb)
        np.testing.assert_allclose(ret[:n, 0], g[f"c{c}_ret"], rtol=1e-12, atol=1e-13)
        np.testing.assert_allclose(adv[:n, 0], g[f"c{c}_adv"], rtol=1e-12, atol=1e-13)
        assert ret[n, 0] == 0 and adv[n, 0] == 0
        only_ret, none = orc.returns(np.append(r, 0.0)[:, None], flag[:, None], gamma)
        assert none is None
        np.testing.assert_allclose(only_ret[:n, 0], g[f"c{c}_ret"], rtol=1e-12, atol=1e-13)
    # chained episodes with a common gamma / lambda: recompute the expectation per episode with the oracle itself on
    # single episodes (pinned above), then check the chained layout reproduces them
    gamma, lamb = 0.97, 0.9
    T = 0
    singles = []
    for c in range(int(g["n_cases"])):
        r, v = g[f"c{c}_rew"], g[f"c{c}_val"]
        n = len(r)
        flag = np.array([0] * (n - 1) + [1, 2], dtype=np.uint8)
        singles.append(orc.returns(np.append(r, 0.0)[:, None], flag[:, None], gamma, values=v[:, None], lamb=lamb))
        rews.append(np.append(r, 0.0)); flags.append(flag); vals.append(v)
        T += n + 1
    R = np.stack([np.concatenate(rews), np.zeros(T)], axis=1)
    F = np.stack([np.concatenate(flags), np.full(T, 4, dtype=np.uint8)], axis=1)
    V = np.stack([np.concatenate(vals), np.ones(T)], axis=1)
    ret, adv = orc.returns(R, F, gamma, values=V, lamb=lamb)
    np.testing.assert_allclose(ret[:, 0], np.concatenate([s[0][:, 0] for s in singles]), rtol=1e-13)
    np.testing.assert_allclose(adv[:, 0], np.concatenate([s[1][:, 0] for s in singles]), rtol=1e-13)
    assert not ret[:, 1].any() and not adv[:, 1].any()
