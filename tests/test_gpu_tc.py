"""
GPU tests of the tensor-core policy path (tcgen05.mma kind::tf32 + TMEM, csrc/sre_policy_tc.cuh): the 3 x TF32 split
must reproduce the fp32 policy (policies/feed_back/fnn.py:138-160) to fp32-level accuracy, in the stand-alone forward
pass and inside the persistent rollout kernel.  The checker is a plain torch fp64 evaluation of the same network.
"""
import numpy as np
import pytest
import torch

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
