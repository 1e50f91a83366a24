"""
GPU tests of the tensor-core policy path (tcgen05.mma kind::tf32 + TMEM, csrc/sre_policy_tc.cuh): the 3 x TF32 split
must reproduce the fp32 policy (policies/feed_back/fnn.py:138-160) to fp32-level accuracy, in the stand-alone forward
pass and inside the persistent rollout kernel.  The checker is a plain torch fp64 evaluation of the same network.
"""
import numpy as np
import pytest
import torch

from conftest import FNN64_CASES, golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


def _weights(sizes, seed=0, scale=0.3):
    rng = np.random.RandomState(seed)
    return (rng.randn(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1))) * scale).astype(np.float32)


def _torch_ref(w, sizes, obs):
    x = torch.as_tensor(obs, dtype=torch.float64)
    off = 0
    for i in range(len(sizes) - 1):
        ni, no = sizes[i], sizes[i + 1]
        W = torch.as_tensor(w[off: off + ni * no].reshape(no, ni), dtype=torch.float64)
        b = torch.as_tensor(w[off + ni * no: off + ni * no + no], dtype=torch.float64)
        off += ni * no + no
        x = x @ W.T + b
        if i < len(sizes) - 2:
            x = torch.tanh(x)
    return x.numpy()


@pytest.mark.parametrize("n_obs,n_act", [(6, 1), (5, 1), (8, 2), (4, 1), (2, 1), (3, 1)])
@pytest.mark.parametrize("n", [1, 127, 128, 1000, 70001])
def test_policy_forward_tensor_core_matches_fp32(E, n_obs, n_act, n):
    from simurlacra_b200 import _lib as L

    sizes = [n_obs, 64, 64, n_act]
    w = _weights(sizes, seed=n_obs)
    wd = torch.from_numpy(w).to(DEV)
    rng = np.random.RandomState(n)
    obs = rng.uniform(-3, 3, size=(n, n_obs)).astype(np.float32)
    obs_d = torch.from_numpy(obs).to(DEV)
    ref = _torch_ref(w, sizes, obs)
    out = {}
    for prec in (L.PREC_FP32, L.PREC_TF32X3, L.PREC_TF32):
        pol = E.make_policy(L.POL_FNN, wd, sizes=sizes, precision=prec)
        out[prec] = E.policy_forward(pol, obs_d, n_act).double().cpu().numpy()
        assert np.isfinite(out[prec]).all()
    scale = max(1.0, np.abs(ref).max())
    err = {k: np.abs(v - ref).max() / scale for k, v in out.items()}
    assert err[L.PREC_FP32] < 5e-6, err
    assert err[L.PREC_TF32X3] < 5e-6, err        # the split reaches the accuracy of the fp32 FMA kernel
    assert err[L.PREC_TF32] < 2e-2, err          # single pass: TF32 accuracy
    assert err[L.PREC_TF32] > err[L.PREC_TF32X3]  # ... and it really is a different arithmetic


def _qq_distr(E):
    from simurlacra_b200 import _lib as L

    names = ["gravity_const", "motor_resistance", "motor_back_emf", "mass_rot_pole", "length_rot_pole",
             "damping_rot_pole", "mass_pend_pole", "length_pend_pole", "damping_pend_pole"]
    nom = E.nominal_params(L.QQ)
    div = dict(gravity_const=10, damping_rot_pole=4, damping_pend_pole=4)
    clip = dict(gravity_const=1e-3, motor_resistance=1e-3, damping_rot_pole=1e-9, damping_pend_pole=1e-9)
    return E.make_distr_table([dict(param_index=i, kind="normal", a=nom[i], b=nom[i] / div.get(k, 5),
                                    clip_lo=clip.get(k, 1e-4)) for i, k in enumerate(names)])


@pytest.mark.parametrize("kind_name,n_obs,n_act", [("qq-su", 6, 1), ("qbb", 8, 2), ("qcp-su", 5, 1), ("pend", 3, 1)])
@pytest.mark.parametrize("n_envs", [1, 130, 1000])
def test_rollout_tensor_core_tracks_fp32_rollout(E, kind_name, n_obs, n_act, n_envs):
    """Same seeds, same weights: the tensor-core rollout follows the fp32-FMA rollout (identical DR draws and initial
    states; trajectories agree to fp32 noise over a short horizon; same rows written, same flags)."""
    from simurlacra_b200 import _lib as L

    kind = E.KIND_BY_NAME[kind_name]
    sizes = [n_obs, 64, 64, n_act]
    wd = torch.from_numpy(_weights(sizes, seed=3, scale=0.2)).to(DEV)
    H = 40
    cfg = E.make_cfg(kind, 1 / 250.0 if kind_name != "qbb" else 0.01, H, dtype=torch.float32, act_norm=True, long_pole=True)
    distr = _qq_distr(E) if kind_name == "qq-su" else None
    res = {}
    for prec in (L.PREC_FP32, L.PREC_TF32X3):
        pol = E.make_policy(L.POL_FNN, wd, sizes=sizes, precision=prec)
        res[prec] = E.rollout(cfg, pol, n_envs, H + 1, device=DEV, distr=distr, seed=11, record=True)
    a, b = res[L.PREC_FP32], res[L.PREC_TF32X3]
    torch.testing.assert_close(a["params"], b["params"], rtol=0, atol=0)
    torch.testing.assert_close(a["state"][0], b["state"][0], rtol=0, atol=0)
    torch.testing.assert_close(a["act"][0], b["act"][0], rtol=1e-4, atol=1e-5)
    same_len = (a["length"] == b["length"]).float().mean()
    assert float(same_len) > 0.98
    m = (a["length"] == b["length"]) & (a["length"] >= 10)
    if int(m.sum()):
        torch.testing.assert_close(a["state"][10][m], b["state"][10][m], rtol=2e-3, atol=2e-3)
        torch.testing.assert_close(a["ret"][m], b["ret"][m], rtol=2e-3, atol=2e-3)
    assert int(b["status"].abs().sum()) == 0


def test_rollout_tensor_core_auto_reset_statistics(E):
    from simurlacra_b200 import _lib as L

    sizes = [6, 64, 64, 1]
    wd = torch.from_numpy(_weights(sizes, seed=0)).to(DEV)
    N, H = 20000, 120
    cfg = E.make_cfg(L.QQ, 1 / 250.0, H, dtype=torch.float32, act_norm=True)
    out = {}
    for prec in (L.PREC_FP32, L.PREC_TF32X3, L.PREC_TF32):
        pol = E.make_policy(L.POL_FNN, wd, sizes=sizes, precision=prec)
        r = E.rollout(cfg, pol, N, H + 1, device=DEV, distr=_qq_distr(E), seed=1, auto_reset=True, record=False)
        out[prec] = r["stats"].cpu().numpy()
    for prec in (L.PREC_TF32X3, L.PREC_TF32):
        assert abs(out[prec][L.STAT_STEPS] - out[L.PREC_FP32][L.STAT_STEPS]) <= 0.01 * N * H
        mean_ret = out[prec][L.STAT_SUM_RET] / out[prec][L.STAT_EPISODES]
        ref_ret = out[L.PREC_FP32][L.STAT_SUM_RET] / out[L.PREC_FP32][L.STAT_EPISODES]
        assert abs(mean_ret - ref_ret) < 0.02 * abs(ref_ret) + 1e-3
    # early exit without auto-reset: every lane finishes, the CTA-wide exit must not hang
    pol = E.make_policy(L.POL_FNN, wd, sizes=sizes, precision=L.PREC_TF32X3)
    r = E.rollout(cfg, pol, 777, H + 1, device=DEV, distr=_qq_distr(E), seed=2, record=False)
    assert int(r["length"].min()) >= 1 and int(r["length"].max()) <= H


# ---- the tensor-core rollout kernel against the REFERENCE and the ORACLE (not against the repo's own FMA path) --------
PRECS = {"fp32_fma": 0, "tf32x3": 1}


@pytest.mark.parametrize("tag", list(FNN64_CASES))
@pytest.mark.parametrize("prec", list(PRECS))
def test_fnn64_closed_loop_golden(E, tag, prec):
    """FNN [do, 64, 64, da] in closed loop on QQube / QCartPole / QBallBalancer / Pendulum: the float32 production build
    (CUDA-core FMA and tcgen05 3 x TF32 hidden layer) against the run of the UNMODIFIED reference (torch float32 policy
    on the CPU, numpy float64 env; tests/golden/fnn64_closed_loop_envs.npz): states, observations and actions within
    1e-4 over the whole episode, episode length equal."""
    g = golden("fnn64_closed_loop_envs")
    kind_name, kw = FNN64_CASES[tag]
    kind = E.KIND_BY_NAME[kind_name]
    sizes = [int(s) for s in g[f"{tag}_sizes"]]
    S, O, A = g[f"{tag}_states"], g[f"{tag}_obs"], g[f"{tag}_acts"]
    H, ms = S.shape[0] - 1, int(g[f"{tag}_max_steps"])
    cfg = E.make_cfg(kind, float(g[f"{tag}_dt"]), ms, dtype=torch.float32, act_norm=bool(g[f"{tag}_actnorm"]), **kw)
    wd = torch.from_numpy(g[f"{tag}_weights"]).to(DEV)
    pol = E.make_policy(E_L().POL_FNN, wd, sizes=sizes, precision=PRECS[prec])
    init = torch.as_tensor(g[f"{tag}_init"][None, :], dtype=torch.float32).to(DEV).contiguous()
    res = E.rollout(cfg, pol, 1, ms + 1, device=DEV, init=init)
    assert int(res["length"][0]) == H, (tag, prec, int(res["length"][0]), H)
    st = res["state"][: H + 1, 0].double().cpu().numpy()
    ob = res["obs"][: H + 1, 0].double().cpu().numpy()
    ac = res["act"][:H, 0].double().cpu().numpy()
    es = (np.abs(st - S) / np.maximum(1.0, np.abs(S))).max()
    eo = (np.abs(ob - O) / np.maximum(1.0, np.abs(O))).max()
    ea = (np.abs(ac - A) / np.maximum(1.0, np.abs(A))).max()
    print(f"fnn64 {tag} {prec}: max rel err state {es:.2e} obs {eo:.2e} act {ea:.2e} over {H} steps")
    assert es < 1e-4 and eo < 1e-4 and ea < 1e-4, (tag, prec, es, eo, ea)
    np.testing.assert_allclose(res["rew"][:H, 0].double().cpu().numpy(), g[f"{tag}_rews"], rtol=1e-4, atol=1e-5)
    assert int(res["status"][0]) == 0


def E_L():
    from simurlacra_b200 import _lib as L

    return L


@pytest.mark.parametrize("kind_name,dt,H,kw", [("qq-su", 1 / 250.0, 600, {}), ("qcp-su", 1 / 250.0, 600, dict(long_pole=True)),
                                               ("qbb", 0.01, 500, {}), ("pend", 1 / 50.0, 400, {})])
@pytest.mark.parametrize("prec", list(PRECS))
def test_rollout_fp32_fnn64_matches_oracle(E, orc, kind_name, dt, H, kw, prec):
    """float32 twin of test_gpu_engine.py::test_rollout_fp64_matches_oracle for the [64, 64] policy: the whole fused path
    of the PRODUCTION build (sampled init states + ActNorm + FNN [do, 64, 64, da] + step + records) against the fp64
    oracle on the same seeded inputs, per step, for both arithmetics of the hidden layer.  Small policy weights keep the
    closed loops in their well-conditioned regime (see test_fp64_sensitivity_of_the_golden_cases for the other one)."""
    L = E_L()
    kind = E.KIND_BY_NAME[kind_name]
    d = E.dims(kind)
    N, Hs = 256, min(H, 200)
    lp = kw.get("long_pole", False)
    sizes = [d["do"], 64, 64, d["da"]]
    w = _weights(sizes, seed=17, scale=0.08)
    cfg = E.make_cfg(kind, dt, Hs, dtype=torch.float32, act_norm=True, **kw)
    pol = E.make_policy(L.POL_FNN, torch.from_numpy(w).to(DEV), sizes=sizes, precision=PRECS[prec])
    res = E.rollout(cfg, pol, N, Hs + 1, device=DEV, seed=23, env_offset=5)
    ref = orc.rollout_batch(kind, dt, Hs, N, Hs, policy=orc.make_policy(orc.POL_FNN, w, sizes=sizes), long_pole=lp,
                            act_norm=True, seed=23, env_offset=5)
    ln = res["length"].cpu().numpy()
    same = ln == ref["length"]
    assert same.mean() >= 0.99 and np.abs(ln - ref["length"]).max() <= 1, (same.mean(), np.abs(ln - ref["length"]).max())
    st, ob, ac, rw = (res[k].double().cpu().numpy() for k in ("state", "obs", "act", "rew"))
    worst = np.zeros(N)
    for n in np.nonzero(same)[0]:
        T = ln[n]
        es = (np.abs(st[: T + 1, n] - ref["state"][: T + 1, n]) / np.maximum(1.0, np.abs(ref["state"][: T + 1, n]))).max()
        eo = (np.abs(ob[: T + 1, n] - ref["obs"][: T + 1, n]) / np.maximum(1.0, np.abs(ref["obs"][: T + 1, n]))).max()
        ea = (np.abs(ac[:T, n] - ref["act"][:T, n]) / np.maximum(1.0, np.abs(ref["act"][:T, n]))).max()
        er = np.abs(rw[:T, n] - ref["rew"][:T, n]).max()
        worst[n] = max(es, eo, ea, er)
    print(f"fp32 fnn64 {kind_name} {prec}: worst lane {worst.max():.2e}, median {np.median(worst):.2e}, same length "
          f"{same.mean():.4f}")
    if kind_name == "qcp-su":
        # The cart-pole's Coulomb friction mu N sign(x_dot) (quanser_cartpole.py:200-209) is DISCONTINUOUS in the state:
        # when x_dot passes through zero within float32 round-off of a step boundary the float32 and fp64 paths apply
        # opposite friction for one step (1 lane of 256 here, 1e-3).  Everything else is held to 1e-4.
        assert np.quantile(worst, 0.98) < 1e-4 and worst.max() < 5e-3, (kind_name, prec, np.quantile(worst, 0.98), worst.max())
    else:
        assert worst.max() < 1e-4, (kind_name, prec, worst.max())
    ok = same & (worst < 1e-4)
    np.testing.assert_allclose(res["ret"].double().cpu().numpy()[ok], ref["ret"][ok], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("kind_name,dt,H,kw,hid", [("qq-su", 1 / 250.0, 120, {}, 64), ("qcp-su", 1 / 250.0, 100, dict(long_pole=True), 64),
                                                   ("qbb", 0.01, 60, {}, 16), ("qq-su", 1 / 250.0, 90, {}, 32)])
@pytest.mark.parametrize("record", [False, True])
def test_specialised_bodies_and_prepared_starts_reproduce_the_generic_kernel(E, monkeypatch, kind_name, dt, H, kw, hid, record):
    """The compile-time specialised step bodies (SRE_MODE_PLAIN / SRE_MODE_PLAIN_REC) and the prepared episode starts
    parked in shared memory are re-arrangements of the generic kernel (SRE_FORCE_GENERIC / SRE_NO_PREP select the
    generic body / direct episode starts at launch time), checked with auto-reset sampling over several episodes per
    lane: prepared starts are bit-identical to direct ones; the specialised bodies agree with the generic one to float32
    round-off."""
    L = E_L()
    kind = E.KIND_BY_NAME[kind_name]
    d = E.dims(kind)
    sizes = [d["do"], hid, hid, d["da"]]
    wd = torch.from_numpy(_weights(sizes, seed=5, scale=0.3)).to(DEV)
    N = 1000
    cfg = E.make_cfg(kind, dt, H, dtype=torch.float32, act_norm=True, **kw)
    distr = _qq_distr(E) if kind_name == "qq-su" else None
    pol = E.make_policy(L.POL_FNN, wd, sizes=sizes)
    out = {}
    for tag, env in (("fast", {}), ("generic", dict(SRE_FORCE_GENERIC="1")), ("noprep", dict(SRE_NO_PREP="1")),
                     ("neither", dict(SRE_FORCE_GENERIC="1", SRE_NO_PREP="1"))):
        for k in ("SRE_FORCE_GENERIC", "SRE_NO_PREP"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        out[tag] = E.rollout(cfg, pol, N, 4 * H, device=DEV, distr=distr, seed=3, auto_reset=True, record=record)
        torch.cuda.synchronize()
    ref = out["neither"]
    assert float(ref["stats"][L.STAT_EPISODES]) > 2 * N  # several episodes per lane: restarts really happen
    keys = ("ret", "length", "params", "status") + (("traj", "row_flag") if record else ())
    # prepared starts vs direct starts: the same body, the same values in the same order -> the same bits (the sums of
    # the statistics vector are accumulated with atomics in launch order: counts exact, sums to fp64 round-off)
    for a, b in (("fast", "noprep"), ("generic", "neither")):
        for k in keys:
            assert torch.equal(out[a][k], out[b][k]), (a, b, k)
        sa, sb = out[a]["stats"].cpu().numpy(), out[b]["stats"].cpu().numpy()
        for i in (L.STAT_EPISODES, L.STAT_SUM_LEN, L.STAT_STEPS, L.STAT_FAILED, L.STAT_MIN_RET, L.STAT_MAX_RET):
            assert sa[i] == sb[i], (a, b, i)
        np.testing.assert_allclose(sa, sb, rtol=1e-12)
    # specialised vs generic body: the same formulas, but the compiler contracts multiply-adds differently in the two
    # bodies, so first-episode values agree to float32 round-off (lengths of a few lanes may move by a step)
    a, b = out["fast"], ref
    assert torch.equal(a["params"], b["params"])
    same = a["length"] == b["length"]
    assert float(same.float().mean()) > 0.99
    torch.testing.assert_close(a["ret"][same], b["ret"][same], rtol=2e-4, atol=2e-4)
    assert abs(float(a["stats"][L.STAT_STEPS]) - float(b["stats"][L.STAT_STEPS])) <= 0.01 * float(b["stats"][L.STAT_STEPS])


@pytest.mark.parametrize("kind_name,sizes", [("QQ", [6, 64, 64, 1]), ("QCP", [5, 64, 64, 1]), ("QBB", [8, 64, 64, 2]),
                                             ("QBB", [8, 16, 16, 2]), ("QQ", [6, 16, 16, 1])])
def test_launch_argument_weights_are_bit_identical(E, kind_name, sizes):
    """sre_policy.weights_host: the small layers of the tensor-core policy (and the whole FNN[16,16]) travel as launch
    arguments and are read as uniform-register operands (rollout_kernel<.., SRE_PV_TCW / SRE_PV_FNN16W>) instead of shared
    memory.  Same arithmetic in the same order: returns, lengths, records and statistics are bit-identical to the launch
    without the host copy, in the plain, the recording and the generic (exploration noise) bodies."""
    from simurlacra_b200 import _lib as L

    kind = getattr(L, kind_name)
    rng = np.random.RandomState(4)
    w = (rng.randn(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(3))) * 0.3).astype(np.float32)
    wd = torch.from_numpy(w).to(DEV)
    cfg = E.make_cfg(kind, 0.01, 80, dtype=torch.float32, act_norm=True, **(dict(long_pole=True) if kind_name == "QCP" else {}))
    N = 1000
    for record, noise in ((False, False), (True, False), (True, True)):
        kw = dict(act_noise=L.ACT_NOISE_NORMAL, act_noise_scale=[0.1] * sizes[-1]) if noise else {}
        out = []
        for host in (None, torch.from_numpy(w)):
            pol = E.make_policy(L.POL_FNN, wd, sizes=sizes, weights_host=host, **kw)
            out.append(E.rollout(cfg, pol, N, 81, device=DEV, seed=6, auto_reset=True, record=record))
        a, b = out
        assert torch.equal(a["ret"], b["ret"]) and torch.equal(a["length"], b["length"])
        # the statistics are accumulated with atomics across CTAs: the sums agree to the order of the additions
        assert torch.allclose(a["stats"], b["stats"], rtol=1e-12, atol=0)
        if record:
            assert torch.equal(a["traj"].view(torch.int32), b["traj"].view(torch.int32)) and torch.equal(a["row_flag"], b["row_flag"])
