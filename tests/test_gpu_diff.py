"""
GPU tests of the differentiable cart-pole path (config C5): forward values, adjoint gradients and per-step
Jacobians against torch autograd run on the unmodified reference (tests/golden/step_diff_qcp.npz), plus an
fp64 finite-difference check at full horizon.
"""
import numpy as np
import pytest
import torch

from conftest import golden
from test_gpu_engine import DEV, _t

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


def _cfg(E, simple, dtype, dt=1 / 20.0):
    from simurlacra_b200 import _lib as L

    return E.make_cfg(L.QCP, dt, 300, dtype=dtype, simple_dynamics=simple, long_pole=True, wild_init=2)


@pytest.mark.parametrize("simple", [False, True])
@pytest.mark.parametrize("tag,dtype,rtol", [("f64", torch.float64, 1e-8), ("f32", torch.float32, 2e-3)])
def test_forward_and_adjoint_match_reference_autograd(E, simple, tag, dtype, rtol):
    g = golden("step_diff_qcp")
    key = f"{'simple' if simple else 'full'}_{tag}"
    s0, acts = g[f"{key}_s0"], g[f"{key}_acts"]
    H, B = acts.shape[0], acts.shape[1]
    cfg = _cfg(E, simple, dtype)
    # wrappers of the reference path: ObsActCat -> ActNorm de-normalise (6 * a for the +-6 V bounds)
    act = _t(6.0 * acts[:, :, 0], dtype)
    states, thdd, rew, done = E.qcp_diff_forward(cfg, None, _t(s0, dtype), act)
    ftol = 1e-9 if dtype == torch.float64 else 2e-4
    np.testing.assert_allclose(states[1:].cpu().numpy(), g[f"{key}_states"], rtol=ftol, atol=ftol)
    np.testing.assert_allclose(rew.cpu().numpy(), g[f"{key}_rews"], rtol=ftol, atol=ftol)
    np.testing.assert_allclose(thdd[1:].cpu().numpy(), g[f"{key}_thdd"], rtol=10 * ftol, atol=10 * ftol)
    if dtype == torch.float64:
        np.testing.assert_array_equal(done.cpu().numpy().astype(bool), g[f"{key}_dones"].astype(bool))
    # loss of the golden run: sum of rewards + 0.1 * sum of the final state
    g_states = torch.zeros(H + 1, B, 4, dtype=dtype, device=DEV)
    g_states[H] = 0.1
    g_rew = torch.ones(H, B, dtype=dtype, device=DEV)
    g_s0, g_act, _ = E.qcp_diff_backward(cfg, None, states, thdd, act, g_states, g_rew)
    ref_s0, ref_a = g[f"{key}_g_s0"], g[f"{key}_g_a"][:, :, 0]
    scale = max(np.abs(ref_s0).max(), 1.0)
    assert np.abs(g_s0.cpu().numpy() - ref_s0).max() < rtol * scale, np.abs(g_s0.cpu().numpy() - ref_s0).max()
    # d/d(normalised action) = 6 * d/d(volts)
    ga = 6.0 * g_act.cpu().numpy()
    assert np.abs(ga - ref_a).max() < rtol * max(np.abs(ref_a).max(), 1.0)
    # actions beyond the clamp have exactly zero gradient through the dynamics, only the reward term remains
    assert np.isfinite(ga).all()


@pytest.mark.parametrize("simple", [False, True])
def test_jacobian_matches_reference_autograd(E, simple):
    g = golden("step_diff_qcp")
    key = f"{'simple' if simple else 'full'}_f64"
    cfg = _cfg(E, simple, torch.float64)
    s = _t(g[f"{key}_s0"][:8], torch.float64)
    a = _t(6.0 * g[f"{key}_acts"][0, :8, 0], torch.float64)
    jac = E.qcp_diff_jacobian(cfg, None, s, a).cpu().numpy()
    ref = g[f"{key}_jac"].copy()
    ref[:, :, 4] /= 6.0  # reference column is d/d(normalised action)
    np.testing.assert_allclose(jac, ref, rtol=1e-8, atol=1e-9)


def test_adjoint_full_horizon_finite_differences(E):
    """Config C5 shape (300 steps, dt = 1/20) in fp64: the adjoint of sum_t r_t agrees with central finite
    differences of the forward kernel for a few coordinates."""
    rng = np.random.RandomState(2)
    B, H = 16, 300
    cfg = _cfg(E, False, torch.float64)
    s0 = rng.uniform(-1, 1, size=(B, 4)) * np.array([0.02, np.pi, 0.0, np.pi / 180])
    acts = rng.uniform(-1, 1, size=(H, B)) * 6.0
    s0_t, a_t = _t(s0, torch.float64), _t(acts, torch.float64)
    states, thdd, rew, _ = E.qcp_diff_forward(cfg, None, s0_t, a_t)
    g_s0, g_act, _ = E.qcp_diff_backward(cfg, None, states, thdd, a_t, None, torch.ones_like(rew))

    def total(s0v, av):
        return E.qcp_diff_forward(cfg, None, _t(s0v, torch.float64), _t(av, torch.float64))[2].sum(0).cpu().numpy()

    eps = 1e-6
    for i in (1, 3):
        p, m = s0.copy(), s0.copy()
        p[:, i] += eps
        m[:, i] -= eps
        fd = (total(p, acts) - total(m, acts)) / (2 * eps)
        ad = g_s0[:, i].cpu().numpy()
        ok = np.abs(fd - ad) <= 1e-4 * np.maximum(1.0, np.abs(ad))
        assert ok.mean() >= 0.8, (i, fd, ad)  # sign(x_dot) / clamp kinks make a few lanes non-differentiable
    for t in (0, 150, 299):
        p, m = acts.copy(), acts.copy()
        p[t] += eps
        m[t] -= eps
        fd = (total(s0, p) - total(s0, m)) / (2 * eps)
        ad = g_act[t].cpu().numpy()
        ok = np.abs(fd - ad) <= 1e-4 * np.maximum(1.0, np.abs(ad))
        assert ok.mean() >= 0.8, (t, fd, ad)


# ---- observation-based step, straight-through clamp, batched differentiable rollout ----
def _wgrad_env(max_steps, simple, dtype):
    from simurlacra_b200.environment_wrappers import ActNormWrapper
    from simurlacra_b200.environment_wrappers.action_normalization import ObsActCatWrapper
    from simurlacra_b200.environments.pysim.quanser_cartpole import QCartPoleSwingUpSim

    sim = QCartPoleSwingUpSim(dt=1 / 20.0, max_steps=max_steps, long=True, simple_dynamics=simple, wild_init=True,
                              device=DEV, dtype=dtype)
    return ObsActCatWrapper(ActNormWrapper(sim))


@pytest.mark.parametrize("simple", [False, True])
def test_step_diff_obs_based_matches_reference_autograd(simple):
    """`step_diff` (quanser_cartpole.py:232-254, :332-340): theta rebuilt from (sin, cos), next OBSERVATION returned;
    values and gradients w.r.t. the observation and the normalised action against the reference's autograd (fp64)."""
    g = golden("diff_extras")
    key = "sd_simple" if simple else "sd_full"
    env = _wgrad_env(300, simple, torch.float64)
    obs = _t(g[f"{key}_obs"], torch.float64).requires_grad_(True)
    act = _t(g[f"{key}_act"], torch.float64).requires_grad_(True)
    nobs, rew, done, info = env.wrapped_env.step_diff(obs, act, _t(g[f"{key}_thdd"], torch.float64))
    np.testing.assert_allclose(nobs.detach().cpu().numpy(), g[f"{key}_nobs"], rtol=1e-9, atol=1e-10)
    np.testing.assert_allclose(rew.detach().cpu().numpy(), g[f"{key}_rew"], rtol=1e-9, atol=1e-11)
    np.testing.assert_array_equal(done.cpu().numpy().astype(bool), g[f"{key}_done"].astype(bool))
    np.testing.assert_allclose(info["th_ddot"].detach().cpu().numpy(), g[f"{key}_thdd_next"], rtol=1e-8, atol=1e-9)
    loss = (nobs * _t(g[f"{key}_wts"], torch.float64)).sum() + rew.sum()
    g_o, g_a = torch.autograd.grad(loss, [obs, act])
    np.testing.assert_allclose(g_o.cpu().numpy(), g[f"{key}_g_obs"], rtol=1e-7, atol=1e-8)
    np.testing.assert_allclose(g_a.cpu().numpy(), g[f"{key}_g_act"], rtol=1e-7, atol=1e-8)


def test_dclamp_is_clamp_forward_identity_backward():
    """DifferentiableClamp / dclamp (action_normalization.py:35-64) against the reference's values and gradient, and
    inside ObsActCatWrapper.step_diff (the carried previous action)."""
    from simurlacra_b200.environment_wrappers.action_normalization import dclamp

    g = golden("diff_extras")
    x = _t(g["dclamp_x"], torch.float64).requires_grad_(True)
    y = dclamp(x, torch.tensor(-1.0, dtype=torch.float64, device=DEV), torch.tensor(1.0, dtype=torch.float64, device=DEV))
    np.testing.assert_allclose(y.detach().cpu().numpy(), g["dclamp_y"], rtol=0, atol=0)
    (gr,) = torch.autograd.grad((y * torch.arange(17.0, dtype=torch.float64, device=DEV)).sum(), [x])
    np.testing.assert_allclose(gr.cpu().numpy(), g["dclamp_g"], rtol=0, atol=0)  # identity: also outside [-1, 1]
    # the wrapper's observation-based step: env stepped with act + previous action, previous action through dclamp
    env = _wgrad_env(300, False, torch.float64)
    B = 8
    th = torch.linspace(-2.0, 2.0, B, dtype=torch.float64, device=DEV)
    prev = torch.linspace(-0.9, 0.9, B, dtype=torch.float64, device=DEV).unsqueeze(1)
    obs = torch.stack([0.01 * th, torch.sin(th), torch.cos(th), 0.1 * th, 0.3 * th], 1)
    act = torch.full((B, 1), 0.4, dtype=torch.float64, device=DEV, requires_grad=True)
    nobs, rew, done, info = env.step_diff(torch.cat([obs, prev], 1), act)
    ref_nobs, ref_rew, _, _ = env.wrapped_env.step_diff(obs, act.detach() + prev)
    np.testing.assert_allclose(nobs.detach().cpu().numpy(), ref_nobs.cpu().numpy(), rtol=1e-12)
    (ga,) = torch.autograd.grad(info["act_prev"].sum(), [act])
    assert torch.all(ga == 1.0) and float(info["act_prev"].max()) <= 1.0  # 0.4 + 0.9 is clamped, its gradient is not


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-8), (torch.float32, 2e-4)])
def test_rollout_tensor_full_matches_reference(dtype, tol):
    """rollout_tensor_full(bsz=4) on ObsActCat(ActNorm(QCartPoleSwingUpSim)) with masked resets against the run of the
    unmodified reference (same policy, the reference's reset states in call order): number, order, lengths and
    completeness of the rollouts, observations / actions / rewards / th_ddot and the per-step Jacobians."""
    from simurlacra_b200.sampling import rollout_tensor_full

    g = golden("diff_extras")
    env = _wgrad_env(int(g["rtf_max_steps_env"]), False, dtype)
    W = torch.as_tensor(g["rtf_W"], dtype=dtype)  # the reference run evaluated the policy in float64

    def policy(obs_to):
        return torch.tanh(obs_to.to(dtype) @ W.T.to(obs_to.device))

    old = torch.get_default_dtype()
    torch.set_default_dtype(dtype)  # rollout_tensor_full hands the policy observations in the default dtype
    try:
        ros = rollout_tensor_full(env, policy, bsz=int(g["rtf_bsz"]), max_steps=int(g["rtf_budget"]),
                                  reset_kwargs=dict(init_state=g["rtf_inits"]))
    finally:
        torch.set_default_dtype(old)
    assert len(ros) == int(g["rtf_n"])
    for i, ro in enumerate(ros):
        ref_obs = g[f"rtf{i}_obs"]
        assert len(ro) == ref_obs.shape[0] - 1, (i, len(ro))
        assert bool(ro.done[-1]) == bool(g[f"rtf{i}_complete"])
        np.testing.assert_allclose(ro.observations, ref_obs, rtol=tol, atol=tol)
        np.testing.assert_allclose(ro.actions, g[f"rtf{i}_act"], rtol=max(tol, 1e-6), atol=max(tol, 1e-6))
        np.testing.assert_allclose(ro.rewards, g[f"rtf{i}_rew"], rtol=tol, atol=tol)
        np.testing.assert_allclose(ro.th_ddot, g[f"rtf{i}_th_ddot"], rtol=10 * tol, atol=10 * tol)
        og, rg = ro.env_infos["obs_grad"], ro.env_infos["rew_grad"]
        assert og.shape == g[f"rtf{i}_obs_grad"].shape and rg.shape == g[f"rtf{i}_rew_grad"].shape
        sc = max(1.0, np.abs(g[f"rtf{i}_obs_grad"]).max())
        assert np.abs(og - g[f"rtf{i}_obs_grad"]).max() < 50 * tol * sc
        assert np.abs(rg - g[f"rtf{i}_rew_grad"]).max() < 50 * tol * max(1.0, np.abs(g[f"rtf{i}_rew_grad"]).max())


def test_rollout_tensor_full_kernel_jacobians_equal_autograd_and_sampler():
    """The one-launch Jacobian route must equal the reference's five-autograd-passes route (forced by hiding the chain
    from the kernel path), and ParallelRolloutSamplerTensor collects at least `min_steps` steps of such rollouts."""
    import simurlacra_b200.sampling.rollout_tensor as RT
    from simurlacra_b200.sampling import ParallelRolloutSamplerTensor, rollout_tensor_full

    env = _wgrad_env(15, False, torch.float64)
    W = torch.tensor([[0.5, -0.4, 0.3, 0.1]], dtype=torch.float64)
    policy = lambda o: torch.tanh(o.double() @ W.T.to(o.device))
    a = rollout_tensor_full(env, policy, bsz=16, max_steps=16 * 25, seed=4)
    orig = RT._kernel_chain
    RT._kernel_chain = lambda e: None
    try:
        b = rollout_tensor_full(env, policy, bsz=16, max_steps=16 * 25, seed=4)
    finally:
        RT._kernel_chain = orig
    assert len(a) == len(b) and len(a) >= 16
    for x, y in zip(a, b):
        np.testing.assert_allclose(x.observations, y.observations, rtol=1e-12, atol=1e-13)
        np.testing.assert_allclose(x.env_infos["obs_grad"], y.env_infos["obs_grad"], rtol=1e-8, atol=1e-9)
        np.testing.assert_allclose(x.env_infos["rew_grad"], y.env_infos["rew_grad"], rtol=1e-8, atol=1e-9)
    assert len({tuple(np.round(r.observations[0], 9)) for r in a}) > 8  # the resets draw different initial states
    smp = ParallelRolloutSamplerTensor(env, policy, min_steps=500, seed=1, bsz=32)
    ros = smp.sample()
    assert sum(len(r) for r in ros) >= 500 and all("obs_grad" in r.env_infos for r in ros)


@pytest.mark.parametrize("simple", [False, True])
def test_packed_pair_kernels_agree_with_the_scalar_kernels(E, simple, monkeypatch):
    """The float forward pass runs two environments per thread on packed f32x2 arithmetic (qcp_diff_forward2_kernel);
    SRE_DIFF_SCALAR=1 forces the one-environment kernel.  Same formulas, so the tape agrees to float round-off over a
    short horizon (the swing-up amplifies differences later)."""
    H, N = 40, 4096
    cfg = _cfg(E, simple, torch.float32)
    g = torch.Generator().manual_seed(5)
    s0 = ((torch.rand(N, 4, generator=g) - 0.5) * torch.tensor([0.4, 6.0, 1.0, 4.0])).to(DEV)
    act = ((torch.rand(H, N, generator=g) - 0.5) * 16.0).to(DEV)
    thdd0 = (torch.rand(N, generator=g) - 0.5).to(DEV)

    def run():
        return E.qcp_diff_forward(cfg, None, s0, act, thdd0=thdd0)

    packed = [x.cpu().numpy() for x in run()]
    monkeypatch.setenv("SRE_DIFF_SCALAR", "1")
    scalar = [x.cpu().numpy() for x in run()]
    for name, a, b, tol in zip(("states", "thdd", "rew", "done"), packed, scalar,
                               (2e-4, 2e-2, 2e-4, 0)):
        if name == "done":
            assert (a != b).mean() < 1e-4
            continue
        scale = max(np.abs(b).max(), 1.0)
        err = np.abs(a.astype(np.float64) - b).max() / scale
        assert err < tol, (name, err)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-12), (torch.float32, 1e-5)])
def test_launch_argument_constants_agree_with_per_env_parameters(E, dtype, tol):
    """Without a parameter tensor the derived cart-pole constants are computed once on the host and passed as launch
    arguments (constant-bank operands, qcp_diff_*_kernel<.., UNI = true>); with a [N, dp] tensor holding the same nominal
    values every lane derives them itself.  Same formulas: forward tape and gradients agree to round-off."""
    from simurlacra_b200 import _lib as L

    H, N = 30, 512
    cfg = _cfg(E, False, dtype)
    g = torch.Generator().manual_seed(9)
    s0 = ((torch.rand(N, 4, generator=g, dtype=torch.float64) - 0.5) * torch.tensor([0.4, 6.0, 1.0, 4.0], dtype=torch.float64)).to(dtype).to(DEV)
    act = ((torch.rand(H, N, generator=g, dtype=torch.float64) - 0.5) * 14.0).to(dtype).to(DEV)
    g_rew = torch.ones(H, N, dtype=dtype, device=DEV)
    params = torch.tensor(E.nominal_params(L.QCP, long_pole=True), dtype=dtype, device=DEV).repeat(N, 1).contiguous()
    out = []
    for p in (None, params):
        states, thdd, rew, done = E.qcp_diff_forward(cfg, p, s0, act)
        g_s0, g_act, _ = E.qcp_diff_backward(cfg, p, states, thdd, act, g_rew=g_rew)
        out.append((states, rew, g_s0, g_act))
    for a, b in zip(*out):
        scale = max(1.0, float(b.abs().max()))
        assert float((a - b).abs().max()) < tol * scale * 50
