"""
GPU parity tests proper: the CUDA engine (through the C-ABI, `simurlacra_b200.engine`) against
  * the golden vectors produced by the unmodified reference (tests/golden),
  * the CPU oracle (oracle/sre_oracle.c) on the same seeded inputs.
Tolerances are the ones north_star states: fp64 1e-10 relative per step and 1e-6 over the first 100 steps;
fp32 1e-4 over the horizon for the non-chaotic envs; done flags and episode lengths bit-exact.
"""
import numpy as np
import pytest
import torch

from conftest import ENV_FILES, OMO_TASK, STAB_TASKS, golden

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module")
def E():
    import simurlacra_b200.engine as eng

    return eng


def _cfg(E, kind_name, dt, max_steps, kw, dtype, **extra):
    from simurlacra_b200 import _lib as L

    kind = E.KIND_BY_NAME[kind_name]
    task = None
    if kind_name == "omo":
        task = E.make_task(kind, state_des=OMO_TASK["state_des"], Q=OMO_TASK["Q"], R=OMO_TASK["R"])
    stab = kw.get("stab")
    if stab:  # QQubeStabSim / QCartPoleStabSim: same kernels, other task (and state space for the cart-pole)
        t = STAB_TASKS[stab]
        task = E.make_task(kind, state_des=t["state_des"], Q=t["Q"], R=t["R"],
                           rew_kind=L.REW_EXP if t["rew"] == "exp" else L.REW_QUADR, radial_idcs=[1],
                           final_flags=(L.FR_ENABLED | L.FR_STATE_DEP | L.FR_TIME_DEP) if t["final"] else 0)
        extra = dict(extra, flags=L.F_QCP_STAB if stab == "qcp-st" else 0)
    if kw.get("disc_act"):
        extra = dict(extra, flags=extra.get("flags", 0) | L.F_DISC_ACT)
    return E.make_cfg(kind, dt, max_steps, dtype=dtype, simple_dynamics=kw.get("simple_dynamics", False),
                      long_pole=kw.get("long_pole", False), task=task, **extra)


def _t(a, dtype):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(DEV).contiguous()


@pytest.mark.parametrize("fname,kind_name,kw", ENV_FILES)
def test_single_step_fp64_matches_reference(E, fname, kind_name, kw):
    """Every golden transition (state_t, hidden_t, act_t) -> state_{t+1}, stepped as one batch: 1e-10 relative."""
    g = golden(fname)
    d = E.dims(E.KIND_BY_NAME[kind_name])
    nhs = max(d["nh"], 1)
    for c in range(int(g["n_cases"])):
        cfg = _cfg(E, kind_name, float(g["dt"]), int(g["max_steps"]), kw, torch.float64)
        S, A, R, D, Hd = g[f"c{c}_states"], g[f"c{c}_actions"], g[f"c{c}_rews"], g[f"c{c}_dones"], g[f"c{c}_hidden"]
        T = len(A)
        params = _t(np.tile(g[f"c{c}_params"], (T, 1)), torch.float64)
        state = _t(S[:T], torch.float64)
        hidden = torch.zeros(T, nhs, dtype=torch.float64, device=DEV)
        if d["nh"]:
            hidden[:, : d["nh"]] = _t(Hd[:T], torch.float64)
        curr = torch.arange(T, dtype=torch.int32, device=DEV)
        obs = torch.empty(T, d["do"], dtype=torch.float64, device=DEV)
        rew = torch.empty(T, dtype=torch.float64, device=DEV)
        done = torch.empty(T, dtype=torch.uint8, device=DEV)
        status = torch.empty(T, dtype=torch.int32, device=DEV)
        E.step(cfg, params, _t(A, torch.float64), state, hidden, curr, obs, rew, done, status)
        ns = state.cpu().numpy()
        scale = np.maximum(1.0, np.abs(S[1: T + 1]))
        assert (np.abs(ns - S[1: T + 1]) / scale).max() < 1e-10
        np.testing.assert_allclose(rew.cpu().numpy(), R, rtol=1e-10, atol=1e-12)
        np.testing.assert_array_equal(done.cpu().numpy().astype(bool), D.astype(bool))
        np.testing.assert_allclose(obs.cpu().numpy(), g[f"c{c}_obs"][1: T + 1], rtol=1e-10, atol=1e-11)
        if d["nh"]:
            np.testing.assert_allclose(hidden[:, : d["nh"]].cpu().numpy(), Hd[1: T + 1], rtol=1e-9, atol=1e-10)
        assert int(status.abs().sum()) == 0
        assert (curr.cpu().numpy() & 0x3FFFFFFF).tolist() == list(range(1, T + 1))  # bit 30: final reward yielded


def _rollout_case(E, g, c, kind_name, kw, dtype):
    from simurlacra_b200 import _lib as L

    d = E.dims(E.KIND_BY_NAME[kind_name])
    H = int(g["max_steps"])
    cfg = _cfg(E, kind_name, float(g["dt"]), H, kw, dtype)
    A = g[f"c{c}_actions"]
    acts = np.zeros((H, 1, d["da"]))
    acts[: len(A), 0] = A
    init = g[f"c{c}_init"][None, :]
    pol = E.make_policy(L.POL_ACTIONS)
    res = E.rollout(cfg, pol, 1, H + 1, device=DEV, params=_t(g[f"c{c}_params"][None, :], dtype), init=_t(init, dtype),
                    actions=_t(acts, dtype))
    return res


@pytest.mark.parametrize("fname,kind_name,kw", ENV_FILES)
def test_trajectory_fp64_matches_reference(E, fname, kind_name, kw):
    from simurlacra_b200 import _lib as L

    g = golden(fname)
    for c in range(int(g["n_cases"])):
        res = _rollout_case(E, g, c, kind_name, kw, torch.float64)
        S, R = g[f"c{c}_states"], g[f"c{c}_rews"]
        Tn = int(g[f"c{c}_length"])
        assert int(res["length"][0]) == Tn  # episode length bit-exact
        flags = res["row_flag"][:, 0].cpu().numpy()
        assert list(flags[:Tn]) == [L.ROW_STEP] * (Tn - 1) + [L.ROW_DONE] and flags[Tn] == L.ROW_TERMINAL
        st = res["state"][: Tn + 1, 0].cpu().numpy()
        err = np.abs(st - S) / np.maximum(1.0, np.abs(S))
        assert err[:101].max() < 1e-6, (fname, c, err[:101].max())
        # beyond 100 steps only the well-conditioned envs are compared tightly (qcp/qq swing-ups amplify 1e-16 errors)
        assert err.max() < 1e-5, (fname, c, err.max())
        np.testing.assert_allclose(res["rew"][:Tn, 0].cpu().numpy(), R, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(res["obs"][: Tn + 1, 0].cpu().numpy(), g[f"c{c}_obs"], rtol=1e-5, atol=1e-6)
        assert abs(float(res["ret"][0]) - R.sum()) < 1e-6 * max(1.0, abs(R.sum()))
        assert int(res["status"][0]) == 0


# fp32 (production build) against the reference's fp64 trajectories: north_star's bar -- 1e-4 over the FULL horizon,
# episode lengths bit-exact -- on every golden file but one.  Round 1 compared the swing-up cases over their first 100
# steps only, calling them chaotic; test_fp64_sensitivity_of_the_golden_cases measured it: a 1e-9 perturbation grows by
# at most 35x (qq-su), 44x (qcp-su), 60x (qcp-su simple), 136x (qcp-st) over the horizon, and the float32 build stays
# within 6e-5 everywhere.  The exception is QQubeStabSim (unstable upright equilibrium, amplification up to 1e3,
# float32 error 3e-4 at 600 steps): 1e-4 is asserted there on the well-conditioned window by the sensitivity test.
FP32_TOL = {f: (1e-4, None) for f, _, _ in ENV_FILES}
FP32_TOL["traj_qq-st"] = (1e-3, None)
SENSITIVITY_FILES = ("traj_qq-su", "traj_qcp-su", "traj_qcp-su_simple", "traj_qq-st", "traj_qcp-st", "traj_qcp-st_simple")
EXACT_LENGTH_FP32 = tuple(f for f, _, _ in ENV_FILES)


@pytest.mark.parametrize("fname,kind_name,kw", ENV_FILES)
def test_trajectory_fp32(E, fname, kind_name, kw):
    g = golden(fname)
    tol, upto = FP32_TOL[fname]
    worst = 0.0
    for c in range(int(g["n_cases"])):
        res = _rollout_case(E, g, c, kind_name, kw, torch.float32)
        S = g[f"c{c}_states"]
        Tn = int(g[f"c{c}_length"])
        n = min(Tn, upto) if upto else Tn
        Tl = int(res["length"][0])
        assert Tl == Tn, (fname, c, Tl, Tn)  # episode lengths bit-exact (north_star)
        m = min(n, Tl)
        st = res["state"][: m + 1, 0].double().cpu().numpy()
        err = np.abs(st - S[: m + 1]) / np.maximum(1.0, np.abs(S[: m + 1]))
        worst = max(worst, err.max())
    assert worst < tol, (fname, worst)


@pytest.mark.parametrize("fname,kind_name,kw", [e for e in ENV_FILES if e[0] in SENSITIVITY_FILES])
def test_fp64_sensitivity_of_the_golden_cases(E, fname, kind_name, kw):
    """Conditioning of the large-action swing-up / stabilisation golden cases, measured on the fp64 build itself: every
    case is run twice in fp64, the second time with the initial state moved by 1e-9 (relative to the state scale);
    g(t) = |delta s(t)| / 1e-9 is the amplification of a perturbation.  A float32 trajectory carries a per-step rounding
    error of ~6e-8 of the state scale, so its distance to the fp64 reference is bounded by ~6e-8 * sum_t g(t): the test
    asserts that the float32 build stays within 1e-4 wherever that bound is below 1e-4 (the float32 error is explained
    by conditioning, not by the kernel) and prints the amplification, which is what justifies the one relaxed entry
    of FP32_TOL (QQubeStabSim)."""
    g = golden(fname)
    amp_max = 0.0
    for c in range(int(g["n_cases"])):
        Tn = int(g[f"c{c}_length"])
        S = g[f"c{c}_states"]
        scale = np.maximum(1.0, np.abs(S).max(0))
        base = _rollout_case(E, g, c, kind_name, kw, torch.float64)
        g2 = {k: g[k] for k in g.files}
        eps = 1e-9
        g2[f"c{c}_init"] = g[f"c{c}_init"] + eps * scale  # all of these envs take a full-state init (ds = 4)
        pert = _rollout_case(E, g2, c, kind_name, kw, torch.float64)
        m = min(Tn, int(pert["length"][0]), int(base["length"][0]))
        ds = (np.abs(pert["state"][: m + 1, 0].cpu().numpy() - base["state"][: m + 1, 0].cpu().numpy()) / scale).max(1)
        gt = ds / eps
        amp_max = max(amp_max, gt.max())
        bound = 6e-8 * np.cumsum(gt)
        f32 = _rollout_case(E, g, c, kind_name, kw, torch.float32)
        m32 = min(m, int(f32["length"][0]))
        e32 = (np.abs(f32["state"][: m32 + 1, 0].double().cpu().numpy() - S[: m32 + 1]) / np.maximum(1.0, np.abs(S[: m32 + 1]))).max(1)
        well = bound[: m32 + 1] < 1e-4
        print(f"{fname} c{c}: max amplification {gt.max():.3g} at step {int(gt.argmax())}, well-conditioned steps "
              f"{int(well.sum())}/{m32 + 1}, fp32 err there {e32[well].max() if well.any() else 0:.2e}, overall {e32.max():.2e}")
        if well.any():
            assert e32[well].max() < 1e-4, (fname, c, e32[well].max())
    print(f"{fname}: max amplification over all cases {amp_max:.3g}")


@pytest.mark.parametrize("fname,kind_name,kw", [("traj_omo", "omo", {}), ("traj_qcp-st", "qcp-su", dict(long_pole=True, stab="qcp-st"))])
def test_final_reward_is_yielded_once_when_stepping_past_done(E, orc, fname, kind_name, kw):
    """FinalRewTask adds its final reward ONCE per episode (tasks/final_reward.py:131-135), also when the caller keeps
    stepping after `done` (rollout(stop_on_done=False), manual stepping): the step kernel carries the flag between
    launches.  Checked against the oracle's env object stepped past done with the same actions."""
    from conftest import OMO_TASK, STAB_TASKS
    from test_oracle_golden import _make

    g = golden(fname)
    c = 2
    kind = E.KIND_BY_NAME[kind_name]
    d = E.dims(kind)
    cfg = _cfg(E, kind_name, float(g["dt"]), 0, kw, torch.float64)  # max_steps = inf: only failures end the episode
    e = _make(orc, kind_name, float(g["dt"]), 0, kw)
    e.set_params(g[f"c{c}_params"])
    A = g[f"c{c}_actions"]
    extra = 30
    acts = np.concatenate([A, np.tile(A[-1:], (extra, 1))])
    e.reset(g[f"c{c}_init"])
    params = _t(g[f"c{c}_params"][None, :], torch.float64)
    state = torch.empty(1, d["ds"], dtype=torch.float64, device=DEV)
    hidden = torch.zeros(1, max(d["nh"], 1), dtype=torch.float64, device=DEV)
    curr = torch.zeros(1, dtype=torch.int32, device=DEV)
    obs = torch.empty(1, d["do"], dtype=torch.float64, device=DEV)
    rew = torch.empty(1, dtype=torch.float64, device=DEV)
    done = torch.empty(1, dtype=torch.uint8, device=DEV)
    E.reset(cfg, params, _t(g[f"c{c}_init"][None, :], torch.float64), state, hidden, curr, obs)
    n_done, big = 0, 0
    for t in range(len(acts)):
        _, r_ref, d_ref = e.step(acts[t].copy())
        E.step(cfg, params, _t(acts[t][None, :], torch.float64), state, hidden, curr, obs, rew, done)
        assert bool(done[0]) == bool(d_ref), t
        assert abs(float(rew[0]) - r_ref) <= 1e-9 * max(1.0, abs(r_ref)), (t, float(rew[0]), r_ref)
        n_done += int(d_ref)
        big += int(abs(r_ref) > 100)
        if not np.all(np.isfinite(e.state)):
            break
    assert n_done >= 2  # the episode was really stepped past its end
