"""
TEST INFRASTRUCTURE ONLY -- generates the golden vectors under tests/golden/ by running the UNMODIFIED
reference (`/root/reference/Pyrado/pyrado`, imported through `oracle/ref_shim.py`).

Run once in the build container (the GPU box has no /root/reference):

    python oracle/gen_golden.py

Conventions:
  1. domain parameters are injected as Python floats (fp64),
  2. actions are float64 arrays in the inner env's units,
  3. QBB plate angles are read from the reference after reset and stored,
  4. the cart-pole `_th_ddot` is stored alongside the state,
  5. files store the state AFTER each step, the reward of each step, `done` and the episode length.
"""
import os
import os.path as osp
import sys
import warnings

import numpy as np

HERE = osp.dirname(osp.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shim  # noqa: E402

ref_shim.install()
warnings.filterwarnings("ignore")

import torch  # noqa: E402
import pyrado  # noqa: E402
from pyrado.environments.pysim.ball_on_beam import BallOnBeamDiscSim, BallOnBeamSim  # noqa: E402
from pyrado.environments.pysim.one_mass_oscillator import OneMassOscillatorSim  # noqa: E402
from pyrado.environments.pysim.pendulum import PendulumSim  # noqa: E402
from pyrado.environments.pysim.quanser_ball_balancer import QBallBalancerSim  # noqa: E402
from pyrado.environments.pysim.quanser_cartpole import QCartPoleStabSim, QCartPoleSwingUpSim  # noqa: E402
from pyrado.environments.pysim.quanser_qube import QQubeStabSim, QQubeSwingUpSim  # noqa: E402
from pyrado.environment_wrappers.action_normalization import ActNormWrapper, ObsActCatWrapper  # noqa: E402
from pyrado.domain_randomization.default_randomizers import (  # noqa: E402
    create_default_randomizer_bob,
    create_default_randomizer_omo,
    create_default_randomizer_pend,
    create_default_randomizer_qbb,
    create_default_randomizer_qcp,
    create_default_randomizer_qq,
)

OUT = osp.join(osp.dirname(HERE), "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def make_env(name, **kw):
    if name == "qq-su":
        return QQubeSwingUpSim(dt=kw.get("dt", 1 / 250.0), max_steps=kw.get("max_steps", 600))
    if name == "qcp-su":
        return QCartPoleSwingUpSim(dt=kw.get("dt", 1 / 250.0), max_steps=kw.get("max_steps", 600),
                                   long=kw.get("long", True), simple_dynamics=kw.get("simple_dynamics", False),
                                   wild_init=kw.get("wild_init", "True"))
    if name == "pend":
        return PendulumSim(dt=kw.get("dt", 1 / 50.0), max_steps=kw.get("max_steps", 400),
                           init_state=kw.get("init_state", np.array([0.3, -0.2])))
    if name == "qq-st":
        return QQubeStabSim(dt=kw.get("dt", 1 / 250.0), max_steps=kw.get("max_steps", 600))
    if name == "qcp-st":
        return QCartPoleStabSim(dt=kw.get("dt", 1 / 250.0), max_steps=kw.get("max_steps", 600),
                                long=kw.get("long", True), simple_dynamics=kw.get("simple_dynamics", False))
    if name == "qbb":
        return QBallBalancerSim(dt=kw.get("dt", 0.01), max_steps=kw.get("max_steps", 500),
                                simple_dynamics=kw.get("simple_dynamics", False))
    if name == "bob":
        return BallOnBeamSim(dt=kw.get("dt", 0.01), max_steps=kw.get("max_steps", 500))
    if name == "bob-d":
        return BallOnBeamDiscSim(dt=kw.get("dt", 0.01), max_steps=kw.get("max_steps", 500))
    if name == "omo":
        return OneMassOscillatorSim(dt=kw.get("dt", 0.02), max_steps=kw.get("max_steps", 300),
                                    task_args=dict(state_des=np.array([0.5, 0.0])))
    raise ValueError(name)


RANDOMIZERS = {
    "qq-su": create_default_randomizer_qq,
    "qcp-su": create_default_randomizer_qcp,
    "pend": create_default_randomizer_pend,
    "qq-st": create_default_randomizer_qq,
    "qcp-st": create_default_randomizer_qcp,
    "qbb": create_default_randomizer_qbb,
    "bob": create_default_randomizer_bob,
    "bob-d": create_default_randomizer_bob,
    "omo": create_default_randomizer_omo,
}


def param_vector(env):
    """Domain parameters in the key order of get_nominal_domain_param()."""
    keys = list(env.get_nominal_domain_param().keys()) if not hasattr(env, "_long") else list(
        env.get_nominal_domain_param(long=env._long).keys())
    dp = env.domain_param
    return keys, np.array([float(dp[k]) for k in keys])


def hidden_of(env):
    if hasattr(env, "_th_ddot") and env._th_ddot is not None:
        return np.array([float(env._th_ddot)])
    if hasattr(env, "plate_angs"):
        return np.array(env.plate_angs, dtype=np.float64).copy()
    return np.zeros(0)


def run_traj(env, init_state, actions, domain_param=None):
    """Hand loop from Pyrado/tests/test_environments.py:87-91 with injected actions."""
    if domain_param is not None:
        domain_param = {k: float(v) for k, v in domain_param.items()}
    env.reset(init_state=np.asarray(init_state, dtype=np.float64).copy(), domain_param=domain_param)
    H = actions.shape[0]
    states = [env.state.copy()]
    obs = [env.observe(env.state)]
    hid = [hidden_of(env)]
    rews, dones = [], []
    length = H
    for t in range(H):
        o, r, d, _ = env.step(actions[t].copy())
        states.append(env.state.copy())
        obs.append(np.asarray(o, dtype=np.float64))
        hid.append(hidden_of(env))
        rews.append(float(r))
        dones.append(bool(d))
        if d:
            length = t + 1
            break
    return dict(states=np.array(states), obs=np.array(obs), hidden=np.array(hid), rews=np.array(rews),
                dones=np.array(dones), length=length)


def gen_env_trajs(name, n_cases=6, H=None, **kw):
    rng = np.random.RandomState(1234 + sum(map(ord, name)))
    env = make_env(name, **kw)
    H = H or env.max_steps
    cases = {}
    keys = None
    for c in range(n_cases):
        env = make_env(name, **kw)
        dp = None
        if c % 2 == 1:
            # randomized parameters, drawn by the reference's own default randomizer, cast to float (fp64)
            pyrado.set_seed(100 + c)
            rz = RANDOMIZERS[name]()
            rz.randomize(1)
            dp = {k: float(v) for k, v in rz.get_params(fmt="dict", dtype="numpy").items()}
        pyrado.set_seed(c)
        if dp is not None:
            env.domain_param = dp
        init = env.init_space.sample_uniform()
        lo, up = env.act_space.bounds
        scale = [0.3, 0.3, 1.5, 1.5, 3.0, 0.05][c % 6]  # includes actions beyond the limits and tiny ones (dead zone)
        actions = rng.uniform(-1, 1, size=(H, lo.size)) * up * scale
        # piecewise-constant segments make the systems move far enough to hit the state bounds
        if c >= 2:
            seg = rng.uniform(-1, 1, size=(H // 25 + 1, lo.size)) * up * scale
            actions = 0.3 * actions + np.repeat(seg, 25, axis=0)[:H]
        tr = run_traj(env, init, actions, dp)
        keys, pv = param_vector(env)
        cases[f"c{c}_init"] = np.asarray(init, dtype=np.float64)
        cases[f"c{c}_params"] = pv
        cases[f"c{c}_actions"] = actions[: tr["length"]]
        for k, v in tr.items():
            cases[f"c{c}_{k}"] = np.asarray(v)
        lo_s, up_s = env.state_space.bounds
        cases[f"c{c}_state_lo"], cases[f"c{c}_state_up"] = lo_s, up_s
        cases[f"c{c}_act_lo"], cases[f"c{c}_act_up"] = lo, up
        print(f"  {name} case {c}: length {tr['length']}/{H} ret {tr['rews'].sum():.6g} randomized={dp is not None}")
    cases["n_cases"] = np.array(n_cases)
    cases["param_keys"] = np.array(keys)
    cases["dt"] = np.array(env.dt)
    cases["max_steps"] = np.array(env.max_steps)
    suffix = "_simple" if kw.get("simple_dynamics") else ""
    np.savez_compressed(osp.join(OUT, f"traj_{name}{suffix}.npz"), **cases)


def gen_fnn_closed_loop():
    """Config C1: QQubeSwingUpSim(dt=1/250, 600 steps), nominal params, FNN [64,64] tanh on the CPU (upstream
    semantics: net(obs); the fork's FNNPolicy hard-codes cart-pole features, so the bare FNN module is used),
    plus the same through ActNormWrapper."""
    from pyrado.policies.feed_back.fnn import FNN

    out = {}
    for tag, hidden, wrap in (("h64", [64, 64], False), ("h16_actnorm", [16, 16], True), ("h32", [32, 32], False)):
        pyrado.set_seed(0)
        env = QQubeSwingUpSim(dt=1 / 250.0, max_steps=600)
        net = FNN(input_size=6, output_size=1, hidden_sizes=hidden, hidden_nonlin=torch.tanh)
        # scale up the output layer so that the closed loop actually moves the pendulum
        with torch.no_grad():
            net.output_layer.weight.mul_(8.0)
        wenv = ActNormWrapper(env) if wrap else env
        init = env.init_space.sample_uniform()
        obs = wenv.reset(init_state=init.copy())
        states, obss, acts, rews = [env.state.copy()], [obs.copy()], [], []
        done = False
        while not done:
            with torch.no_grad():
                a = net(torch.from_numpy(obs).to(torch.get_default_dtype())).numpy().astype(np.float64)
            obs, r, done, _ = wenv.step(a)
            states.append(env.state.copy()); obss.append(obs.copy()); acts.append(a); rews.append(r)
        params = []
        for layer in list(net.hidden_layers) + [net.output_layer]:
            params += [layer.weight.detach().numpy().reshape(-1), layer.bias.detach().numpy().reshape(-1)]
        out[f"{tag}_weights"] = np.concatenate(params).astype(np.float32)
        out[f"{tag}_sizes"] = np.array([6] + hidden + [1])
        out[f"{tag}_init"] = init
        out[f"{tag}_states"] = np.array(states)
        out[f"{tag}_obs"] = np.array(obss)
        out[f"{tag}_acts"] = np.array(acts)
        out[f"{tag}_rews"] = np.array(rews)
        out[f"{tag}_actnorm"] = np.array(wrap)
        print(f"  fnn {tag}: length {len(rews)} ret {np.sum(rews):.6g} max|a| {np.abs(acts).max():.3g}")
    np.savez_compressed(osp.join(OUT, "fnn_closed_loop_qq.npz"), **out)


def gen_policy_forward():
    """FNN / Linear forward known answers (float32 torch on CPU): fnn.py:138-160, linear.py:95-110, features.py."""
    from pyrado.policies.feed_back.fnn import FNN
    from pyrado.policies.features import (FeatureStack, const_feat, identity_feat, squared_feat, sin_feat, cos_feat,
                                          abs_feat, sign_feat, cubic_feat, sinsin_feat, sincos_feat)

    out = {}
    torch.manual_seed(7)
    for tag, (n_in, hidden, n_out) in dict(a=(6, [64, 64], 1), b=(8, [16, 16], 2), c=(5, [32, 32], 1), d=(4, [16], 1)).items():
        net = FNN(input_size=n_in, output_size=n_out, hidden_sizes=hidden, hidden_nonlin=torch.tanh)
        x = torch.randn(32, n_in) * 2.0
        with torch.no_grad():
            y = net(x)
        params = []
        for layer in list(net.hidden_layers) + [net.output_layer]:
            params += [layer.weight.detach().numpy().reshape(-1), layer.bias.detach().numpy().reshape(-1)]
        out[f"fnn_{tag}_weights"] = np.concatenate(params).astype(np.float32)
        out[f"fnn_{tag}_sizes"] = np.array([n_in] + hidden + [n_out])
        out[f"fnn_{tag}_x"] = x.numpy()
        out[f"fnn_{tag}_y"] = y.numpy()
    feats = [const_feat, identity_feat, sign_feat, abs_feat, squared_feat, cubic_feat, sin_feat, cos_feat, sinsin_feat,
             sincos_feat]
    names = ["const", "identity", "sign", "abs", "squared", "cubic", "sin", "cos", "sinsin", "sincos"]
    fs = FeatureStack(*feats)
    x = torch.randn(16, 4)
    W = torch.randn(2, fs.get_num_feat(4)) * 0.3
    with torch.no_grad():
        y = torch.stack([W @ fs(x[i]) for i in range(x.shape[0])])
    out["lin_feats"] = np.array(names)
    out["lin_W"] = W.numpy()
    out["lin_x"] = x.numpy()
    out["lin_y"] = y.numpy()
    # parametrised features of the reference (features.py:157-161, :181-445): feature values and W @ feats for 1-dim inputs
    import functools
    from pyrado.policies.features import ATan2Feat, MultFeat, RBFFeat, RFFeat, bell_feat, sig_feat

    torch.manual_seed(11)
    d = 4
    xp = torch.randn(16, d)
    rf = RFFeat(inp_dim=d, num_feat_per_dim=12, bandwidth=np.array([1.0, 2.0, 0.5, 1.5]))
    rbf_s = RBFFeat(num_feat_per_dim=5, bounds=(np.array([-2.0, -1.0, -3.0, -2.0]), np.array([2.0, 1.0, 3.0, 2.0])))
    rbf_j = RBFFeat(num_feat_per_dim=4, bounds=(np.array([-2.0, -1.0, -3.0, -2.0]), np.array([2.0, 1.0, 3.0, 2.0])), scale=0.7,
                    state_wise_norm=False)
    stacks = dict(
        elem=[const_feat, functools.partial(sig_feat, scale=1.7), functools.partial(bell_feat, scale=0.6), sig_feat,
              MultFeat((0, 2)), MultFeat((1, 2, 3)), ATan2Feat(0, 1), identity_feat],
        rf=[rf], rbf_state=[rbf_s], rbf_joint=[rbf_j])
    for tag, fl in stacks.items():
        vals = []
        for i in range(xp.shape[0]):
            vals.append(torch.cat([f(xp[i]).reshape(-1) for f in fl]))  # 1-dim input like a policy call (RF / RBF give [1, n])
        F = torch.stack(vals)
        Wp = torch.randn(2, F.shape[1]) * 0.3
        out[f"pf_{tag}_feats"] = F.numpy()
        out[f"pf_{tag}_W"] = Wp.numpy()
        out[f"pf_{tag}_y"] = (F @ Wp.T).numpy()
    out["pf_x"] = xp.numpy()
    out["pf_rf_freq"], out["pf_rf_shift"], out["pf_rf_scale"] = rf.freq.numpy(), rf.shift.numpy(), np.float32(rf.scale)
    out["pf_rbf_state_centers"], out["pf_rbf_state_scale"] = rbf_s.centers.numpy(), rbf_s.scale.numpy()
    out["pf_rbf_joint_centers"], out["pf_rbf_joint_scale"] = rbf_j.centers.numpy(), np.float32(rbf_j.scale)
    np.savez_compressed(osp.join(OUT, "policy_forward.npz"), **out)


def gen_qbb_kin():
    env = QBallBalancerSim(dt=0.01, max_steps=500)
    rows = []
    rng = np.random.RandomState(3)
    for i in range(40):
        r_arm = 0.0254 * (1 + 0.3 * rng.uniform(-1, 1)) if i % 2 else 0.0254
        l_plate = 0.275 * (1 + 0.3 * rng.uniform(-1, 1)) if i % 2 else 0.275
        th = rng.uniform(-np.pi / 2.1, np.pi / 2.1) if i > 0 else 0.0
        env.domain_param = dict(arm_radius=float(r_arm), plate_length=float(l_plate))
        ang = env._kin(th)
        rows.append([r_arm, l_plate, th, ang])
    np.savez_compressed(osp.join(OUT, "qbb_kin.npz"), rows=np.array(rows))
    print("  qbb kin: 40 rows")


def gen_step_diff():
    """Config C5 pieces: step_diff_state values and autograd gradients through short horizons (fp64 and fp32)."""
    out = {}
    for simple in (False, True):
        for dtype, tag in ((torch.float64, "f64"), (torch.float32, "f32")):
            torch.set_default_dtype(dtype)
            env = ObsActCatWrapper(ActNormWrapper(QCartPoleSwingUpSim(dt=1 / 20.0, max_steps=300, long=True,
                                                                      simple_dynamics=simple, wild_init=True)))
            inner = env._wrapped_env._wrapped_env
            rng = np.random.RandomState(11)
            B, H = 64, 20
            s0 = rng.uniform(-1, 1, size=(B, 4)) * np.array([0.2, np.pi, 0.5, 2.0])
            acts = rng.uniform(-1.2, 1.2, size=(H, B, 1))  # normalised; some beyond +-1 (clamped, zero gradient)
            s0_t = torch.tensor(s0, dtype=dtype, requires_grad=True)
            a_t = torch.tensor(acts, dtype=dtype, requires_grad=True)
            # the reference keeps bounds / Q / R as float32 tensors; move them to the working dtype
            inner._state_space_bound_lo_tensor = inner._state_space_bound_lo_tensor.to(dtype)
            inner._state_space_bound_up_tensor = inner._state_space_bound_up_tensor.to(dtype)
            inner.state_des, inner.Q, inner.R = inner.state_des.to(dtype), inner.Q.to(dtype), inner.R.to(dtype)
            s = s0_t
            thdd = torch.zeros(B, dtype=dtype)
            states, rews, dones, thdds = [], [], [], []
            for t in range(H):
                s, r, d, info = env.step_diff_state(s, a_t[t], thdd)
                thdd = info["th_ddot"]
                states.append(s); rews.append(r); dones.append(d); thdds.append(thdd)
            total = torch.stack(rews).sum() + 0.1 * s.sum()
            g_s0, g_a = torch.autograd.grad(total, [s0_t, a_t])
            key = f"{'simple' if simple else 'full'}_{tag}"
            out[f"{key}_s0"] = s0
            out[f"{key}_acts"] = acts
            out[f"{key}_states"] = torch.stack(states).detach().numpy()
            out[f"{key}_rews"] = torch.stack(rews).detach().numpy()
            out[f"{key}_dones"] = torch.stack(dones).numpy()
            out[f"{key}_thdd"] = torch.stack(thdds).detach().numpy()
            out[f"{key}_g_s0"] = g_s0.numpy()
            out[f"{key}_g_a"] = g_a.numpy()
            # one-step Jacobians d(next_state, rew)/d(state, act) for the first 8 envs
            s1 = torch.tensor(s0[:8], dtype=dtype, requires_grad=True)
            a1 = torch.tensor(acts[0, :8], dtype=dtype, requires_grad=True)
            ns, r, d, info = env.step_diff_state(s1, a1, torch.zeros(8, dtype=dtype))
            jac = np.zeros((8, 5, 5))
            for i in range(4):
                gs, ga = torch.autograd.grad(ns[:, i].sum(), [s1, a1], retain_graph=True)
                jac[:, i, :4], jac[:, i, 4] = gs.numpy(), ga.numpy()[:, 0]
            gs, ga = torch.autograd.grad(r.sum(), [s1, a1])
            jac[:, 4, :4], jac[:, 4, 4] = gs.numpy(), ga.numpy()[:, 0]
            out[f"{key}_jac"] = jac
            print(f"  step_diff {key}: ret {float(torch.stack(rews).sum()):.6g} |g_s0| {np.abs(g_s0.numpy()).max():.4g}")
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(osp.join(OUT, "step_diff_qcp.npz"), **out)


def gen_dr_moments():
    """Reference DR draws (torch MT stream, float32): only moments/clip behaviour can be compared with the Philox
    engine; the raw draws are stored to pin the dtype/clipping conventions (domain_parameter.py:104-132)."""
    out = {}
    for name, fn in RANDOMIZERS.items():
        pyrado.set_seed(0)
        rz = fn()
        rz.randomize(2000)
        d = rz.get_params(-1, fmt="dict", dtype="numpy")
        out[f"{name}_keys"] = np.array(list(d.keys()))
        out[f"{name}_mean"] = np.array([np.mean(np.array(v, dtype=np.float64)) for v in d.values()])
        out[f"{name}_std"] = np.array([np.std(np.array(v, dtype=np.float64)) for v in d.values()])
        out[f"{name}_min"] = np.array([np.min(np.array(v, dtype=np.float64)) for v in d.values()])
        out[f"{name}_max"] = np.array([np.max(np.array(v, dtype=np.float64)) for v in d.values()])
        out[f"{name}_dtype"] = np.array(str(np.asarray(list(d.values())[0][0]).dtype))
        # distribution table as the reference defines it
        rows = []
        for dp in rz.domain_params:
            kind = type(dp).__name__
            if kind == "NormalDomainParam":
                rows.append([0, dp.mean, dp.std, 0.0, dp.clip_lo, dp.clip_up])
            elif kind == "UniformDomainParam":
                rows.append([1, dp.mean - dp.halfspan, dp.mean + dp.halfspan, 0.0, dp.clip_lo, dp.clip_up])
        out[f"{name}_table"] = np.array(rows, dtype=np.float64)
    np.savez_compressed(osp.join(OUT, "dr_reference.npz"), **out)
    print("  dr moments written")


def gen_returns():
    """discounted_value / gae_returns of the unmodified reference (sampling/step_sequence.py:1049-1116) on random
    complete rollouts (done only at the last step, as rollout() produces them)."""
    from pyrado.sampling.step_sequence import StepSequence, discounted_value, gae_returns

    rng = np.random.RandomState(77)
    out = {}
    cases = [(1, 0.99, 0.95), (2, 0.9, 0.5), (37, 0.99, 0.95), (200, 0.995, 0.97), (64, 1.0, 1.0), (50, 0.0, 0.3)]
    for c, (n, gamma, lamb) in enumerate(cases):
        rew = rng.randn(n)
        val = rng.randn(n + 1)
        ro = StepSequence(rewards=rew, observations=rng.randn(n + 1, 3), actions=rng.randn(n, 1), values=val,
                          complete=True)
        out[f"c{c}_rew"], out[f"c{c}_val"] = rew, val
        out[f"c{c}_gamma_lamb"] = np.array([gamma, lamb])
        out[f"c{c}_ret"] = np.asarray(discounted_value(ro, gamma), dtype=np.float64).copy()
        out[f"c{c}_adv"] = np.asarray(gae_returns(ro, gamma, lamb), dtype=np.float64).copy()
        print(f"  returns case {c}: n={n} G0={out[f'c{c}_ret'][0]:.6g} A0={out[f'c{c}_adv'][0]:.6g}")
    out["n_cases"] = np.array(len(cases))
    # discounted_values over SEVERAL rollouts (step_sequence.py:1075-1090): one 1-D tensor for 'torch', one row per
    # rollout for 'numpy'
    from pyrado.sampling.step_sequence import discounted_values

    for tag, lens in (("eq", (5, 5, 5)), ("ragged", (4, 7, 3))):
        ros = [StepSequence(rewards=rng.randn(n), observations=rng.randn(n + 1, 2), actions=rng.randn(n, 1), complete=True)
               for n in lens]
        out[f"dvs_{tag}_lens"] = np.array(lens)
        out[f"dvs_{tag}_rew"] = np.concatenate([np.asarray(ro.rewards, dtype=np.float64) for ro in ros])
        out[f"dvs_{tag}_torch"] = discounted_values(ros, 0.97, "torch").numpy().astype(np.float64)
        if tag == "eq":
            out[f"dvs_{tag}_numpy"] = np.asarray(discounted_values(ros, 0.97, "numpy"), dtype=np.float64)
        print(f"  discounted_values {tag}: torch shape {out[f'dvs_{tag}_torch'].shape}")
    np.savez_compressed(osp.join(OUT, "returns_reference.npz"), **out)


def gen_stab_envs():
    """the stabilisation variants (other init / state spaces, FinalRewTask on the cart-pole)."""
    gen_env_trajs("qq-st", n_cases=4)
    gen_env_trajs("qcp-st", n_cases=4)
    gen_env_trajs("qcp-st", n_cases=4, simple_dynamics=True)


def gen_wrappers():
    """SURVEY.md 8(f) rank 2, the stateful / shape-changing step-path wrappers: ActDelayWrapper, DownsamplingWrapper,
    ObsPartialWrapper, ObsVelFiltWrapper, StateAugmentationWrapper of the unmodified reference around
    QQubeSwingUpSim(dt=1/250) with injected actions; observations as the wrapper returns them (reset + every step)."""
    from pyrado.environment_wrappers.action_delay import ActDelayWrapper
    from pyrado.environment_wrappers.downsampling import DownsamplingWrapper
    from pyrado.environment_wrappers.observation_partial import ObsPartialWrapper
    from pyrado.environment_wrappers.observation_velfilter import ObsVelFiltWrapper
    from pyrado.environment_wrappers.state_augmentation import StateAugmentationWrapper

    rng = np.random.RandomState(77)
    H = 60
    act = rng.uniform(-3.0, 3.0, size=(H, 1))
    s0 = np.array([0.1, -0.2, 0.3, -0.4])
    dp = dict(mass_pend_pole=0.03, length_pend_pole=0.15)
    cases = dict(
        delay3=lambda e: ActDelayWrapper(e, delay=3),
        delay3_norm=lambda e: ActNormWrapper(ActDelayWrapper(e, delay=3)),
        down4=lambda e: DownsamplingWrapper(e, factor=4),
        partial_mask=lambda e: ObsPartialWrapper(e, mask=[0, 0, 0, 0, 1, 1]),
        partial_keep=lambda e: ObsPartialWrapper(e, idcs=["sin_theta", "cos_theta", "theta_dot"], keep_selected=True),
        velfilt=lambda e: ObsVelFiltWrapper(e, idcs_pos=["theta", "alpha"], idcs_vel=["theta_dot", "alpha_dot"]),
        augment=lambda e: StateAugmentationWrapper(e, domain_param=["mass_pend_pole", "length_pend_pole"]),
        augment_fixed=lambda e: StateAugmentationWrapper(e, domain_param=["mass_pend_pole", "length_pend_pole"], fixed=True),
    )
    out = dict(act=act, init_state=s0, dp_keys=np.array(list(dp.keys())), dp_vals=np.array(list(dp.values())))
    for name, mk in cases.items():
        env = mk(QQubeSwingUpSim(dt=1 / 250.0, max_steps=H))
        a = act / 5.0 if name.endswith("_norm") else act  # normalised units for the ActNorm chain
        obs = [np.asarray(env.reset(init_state=s0.copy(), domain_param=dict(dp)), dtype=np.float64).copy()]
        states, rews = [env.state.copy()], []
        for t in range(H):
            o, r, d, _ = env.step(a[t].copy())
            obs.append(np.asarray(o, dtype=np.float64).copy())
            states.append(env.state.copy())
            rews.append(float(r))
            if d:
                break
        out[name + "_len"] = np.array(len(rews))
        out[name + "_obs"], out[name + "_states"], out[name + "_rews"] = np.array(obs), np.array(states), np.array(rews)
        out[name + "_obs_lo"], out[name + "_obs_up"] = env.obs_space.bound_lo, env.obs_space.bound_up
    np.savez_compressed(osp.join(OUT, "wrappers_reference.npz"), **out)
    print("wrappers_reference.npz", {k: v.shape for k, v in out.items() if k.endswith("_obs")})


def gen_calm_trajs():
    """north_star's fp32 bar (1e-4 over the FULL 600-step horizon, lengths bit-exact) needs cases where the swing-up
    pendulums are NOT chaotic: small voltages (|a| <= 0.5 V) about the hanging equilibrium, nominal and randomized
    parameters.  Same file format as gen_env_trajs."""
    for name, kw in (("qq-su", {}), ("qcp-su", dict(wild_init="False")), ("qcp-su", dict(wild_init="False", simple_dynamics=True))):
        rng = np.random.RandomState(4321 + sum(map(ord, name)) + int(bool(kw.get("simple_dynamics"))))
        cases, keys = {}, None
        n_cases = 4
        for c in range(n_cases):
            env = make_env(name, **kw)
            H = env.max_steps
            dp = None
            if c % 2 == 1:
                pyrado.set_seed(200 + c)
                rz = RANDOMIZERS[name]()
                rz.randomize(1)
                dp = {k: float(v) for k, v in rz.get_params(fmt="dict", dtype="numpy").items()}
            pyrado.set_seed(50 + c)
            if dp is not None:
                env.domain_param = dp
            init = env.init_space.sample_uniform()  # qq: +-[2, 1, .5, .5] deg about hanging; qcp 'False' box
            amp = [0.5, 0.3, 0.5, 0.1][c]
            actions = rng.uniform(-1, 1, size=(H, 1)) * amp
            if c >= 2:  # slow piecewise-constant component
                seg = rng.uniform(-1, 1, size=(H // 50 + 1, 1)) * amp
                actions = 0.4 * actions + 0.6 * np.repeat(seg, 50, axis=0)[:H]
            tr = run_traj(env, init, actions, dp)
            keys, pv = param_vector(env)
            cases[f"c{c}_init"] = np.asarray(init, dtype=np.float64)
            cases[f"c{c}_params"] = pv
            cases[f"c{c}_actions"] = actions[: tr["length"]]
            for k, v in tr.items():
                cases[f"c{c}_{k}"] = np.asarray(v)
            lo_s, up_s = env.state_space.bounds
            lo, up = env.act_space.bounds
            cases[f"c{c}_state_lo"], cases[f"c{c}_state_up"] = lo_s, up_s
            cases[f"c{c}_act_lo"], cases[f"c{c}_act_up"] = lo, up
            print(f"  {name} calm case {c}: length {tr['length']}/{H} ret {tr['rews'].sum():.6g} max|s| "
                  f"{np.abs(tr['states']).max(0)} randomized={dp is not None}")
        cases["n_cases"] = np.array(n_cases)
        cases["param_keys"] = np.array(keys)
        cases["dt"] = np.array(env.dt)
        cases["max_steps"] = np.array(env.max_steps)
        suffix = "_simple" if kw.get("simple_dynamics") else ""
        np.savez_compressed(osp.join(OUT, f"traj_{name}_calm{suffix}.npz"), **cases)


def gen_fnn_closed_loop_envs():
    """FNN [do, 64, 64, da] tanh in closed loop on the reference (torch float32 policy on the CPU, numpy float64 env)
    for QQube, QCartPole, QBallBalancer and Pendulum: the golden cases of the tensor-core policy path.  The output
    layer keeps its default (small) initialisation scaled so that the systems move but stay well conditioned."""
    from pyrado.policies.feed_back.fnn import FNN

    out = {}
    specs = (("qq", lambda: QQubeSwingUpSim(dt=1 / 250.0, max_steps=600), False, 2.0),
             ("qq_norm", lambda: QQubeSwingUpSim(dt=1 / 250.0, max_steps=600), True, 0.25),
             ("qcp", lambda: QCartPoleSwingUpSim(dt=1 / 250.0, max_steps=600, long=True, wild_init="False"), True, 25.0),
             ("qbb", lambda: QBallBalancerSim(dt=0.01, max_steps=500), True, 0.5),
             ("pend", lambda: PendulumSim(dt=1 / 50.0, max_steps=400, init_state=np.array([0.3, -0.2])), False, 1.0))
    for i, (tag, mk, wrap, gain) in enumerate(specs):
        pyrado.set_seed(10 + i)
        env = mk()
        n_in, n_out = env.obs_space.flat_dim, env.act_space.flat_dim
        net = FNN(input_size=n_in, output_size=n_out, hidden_sizes=[64, 64], hidden_nonlin=torch.tanh)
        with torch.no_grad():
            net.output_layer.weight.mul_(gain)
        wenv = ActNormWrapper(env) if wrap else env
        init = env.init_space.sample_uniform()
        o = wenv.reset(init_state=init.copy())
        obs = np.asarray(env.observe(env.state), dtype=np.float64) if tag.startswith("qcp") else np.asarray(o, dtype=np.float64)
        states, obss, acts, rews, hid = [env.state.copy()], [obs.copy()], [], [], [hidden_of(env)]
        done = False
        while not done:
            with torch.no_grad():
                a = net(torch.from_numpy(obs).to(torch.get_default_dtype())).numpy().astype(np.float64)
            o, r, done, _ = wenv.step(a)
            obs = np.asarray(o, dtype=np.float64)
            states.append(env.state.copy()); obss.append(obs.copy()); acts.append(a); rews.append(r); hid.append(hidden_of(env))
        params = []
        for layer in list(net.hidden_layers) + [net.output_layer]:
            params += [layer.weight.detach().numpy().reshape(-1), layer.bias.detach().numpy().reshape(-1)]
        out[f"{tag}_weights"] = np.concatenate(params).astype(np.float32)
        out[f"{tag}_sizes"] = np.array([n_in, 64, 64, n_out])
        out[f"{tag}_init"] = np.asarray(init, dtype=np.float64)
        out[f"{tag}_states"] = np.array(states)
        out[f"{tag}_hidden"] = np.array(hid)
        out[f"{tag}_obs"] = np.array(obss)
        out[f"{tag}_acts"] = np.array(acts)
        out[f"{tag}_rews"] = np.array(rews)
        out[f"{tag}_actnorm"] = np.array(wrap)
        out[f"{tag}_dt"] = np.array(env.dt)
        out[f"{tag}_max_steps"] = np.array(env.max_steps)
        print(f"  fnn64 {tag}: length {len(rews)} ret {np.sum(rews):.6g} max|a| {np.abs(acts).max():.3g} "
              f"max|s| {np.abs(np.array(states)).max(0)}")
    out["tags"] = np.array([s[0] for s in specs])
    np.savez_compressed(osp.join(OUT, "fnn64_closed_loop_envs.npz"), **out)


def gen_diff_extras():
    """the observation-based differentiable step (`step_diff`, theta rebuilt from
    (sin, cos)), the straight-through clamp, and the batched differentiable rollout `rollout_tensor_full` with masked
    resets and per-step Jacobians -- all run on the unmodified reference in float64."""
    from pyrado.environment_wrappers.action_normalization import dclamp
    from pyrado.sampling.rollout import rollout_tensor_full

    out = {}
    dtype = torch.float64
    torch.set_default_dtype(dtype)

    def mk(max_steps, simple):
        env = ObsActCatWrapper(ActNormWrapper(QCartPoleSwingUpSim(dt=1 / 20.0, max_steps=max_steps, long=True,
                                                                  simple_dynamics=simple, wild_init=True)))
        inner = env._wrapped_env._wrapped_env
        inner._state_space_bound_lo_tensor = inner._state_space_bound_lo_tensor.to(dtype)
        inner._state_space_bound_up_tensor = inner._state_space_bound_up_tensor.to(dtype)
        inner.state_des, inner.Q, inner.R = inner.state_des.to(dtype), inner.Q.to(dtype), inner.R.to(dtype)
        return env, inner

    # ---- step_diff (observation based), through ActNormWrapper
    rng = np.random.RandomState(5)
    for simple in (False, True):
        env, inner = mk(300, simple)
        B = 48
        th = rng.uniform(-np.pi, np.pi, size=B)
        obs = np.stack([rng.uniform(-0.2, 0.2, B), np.sin(th), np.cos(th), rng.uniform(-0.5, 0.5, B), rng.uniform(-2, 2, B)], 1)
        act = rng.uniform(-1.2, 1.2, size=(B, 1))
        thdd = rng.uniform(-1, 1, size=B)
        o_t = torch.tensor(obs, requires_grad=True)
        a_t = torch.tensor(act, requires_grad=True)
        nobs, rew, done, info = env._wrapped_env.step_diff(o_t, a_t, torch.tensor(thdd))
        wts = torch.tensor(rng.uniform(-1, 1, size=(5,)))
        loss = (nobs * wts).sum() + rew.sum()
        g_o, g_a = torch.autograd.grad(loss, [o_t, a_t])
        key = "sd_simple" if simple else "sd_full"
        out.update({f"{key}_obs": obs, f"{key}_act": act, f"{key}_thdd": thdd, f"{key}_wts": wts.numpy(),
                    f"{key}_nobs": nobs.detach().numpy(), f"{key}_rew": rew.detach().numpy(), f"{key}_done": done.numpy(),
                    f"{key}_thdd_next": info["th_ddot"].detach().numpy(), f"{key}_g_obs": g_o.numpy(), f"{key}_g_act": g_a.numpy()})
        print(f"  step_diff {key}: |g_obs| {np.abs(g_o.numpy()).max():.4g} done {int(done.sum())}")

    # ---- dclamp: clamp forward, identity backward
    x = torch.linspace(-2.0, 2.0, 17, requires_grad=True)
    y = dclamp(x, torch.tensor(-1.0), torch.tensor(1.0))
    (g,) = torch.autograd.grad((y * torch.arange(17.0)).sum(), [x])
    out.update(dclamp_x=x.detach().numpy(), dclamp_y=y.detach().numpy(), dclamp_g=g.numpy())

    # ---- rollout_tensor_full(bsz=4): the init states the reference's resets draw are logged in call order
    env, inner = mk(12, False)
    W = np.array([[0.8, -0.3, 0.5, 0.2]])

    def policy(obs_to):
        return torch.tanh(obs_to @ torch.tensor(W).T.to(obs_to))

    inits = []
    orig_reset = env.reset

    def logging_reset(*a, **k):
        o = orig_reset(*a, **k)
        inits.append(np.array(o, dtype=np.float64).copy())
        return o

    env.reset = logging_reset
    pyrado.set_seed(3)
    ros = rollout_tensor_full(env, policy, bsz=4, max_steps=4 * 20)
    out["rtf_W"], out["rtf_inits"], out["rtf_n"] = W, np.array(inits), np.array(len(ros))
    out["rtf_max_steps_env"], out["rtf_budget"], out["rtf_bsz"] = np.array(12), np.array(80), np.array(4)
    for i, ro in enumerate(ros):
        out[f"rtf{i}_obs"] = np.asarray(ro.observations, dtype=np.float64)
        out[f"rtf{i}_act"] = np.asarray(ro.actions, dtype=np.float64)
        out[f"rtf{i}_rew"] = np.asarray(ro.rewards, dtype=np.float64)
        out[f"rtf{i}_obs_grad"] = np.asarray(ro.env_infos["obs_grad"], dtype=np.float64)
        out[f"rtf{i}_rew_grad"] = np.asarray(ro.env_infos["rew_grad"], dtype=np.float64)
        out[f"rtf{i}_th_ddot"] = np.asarray(ro.th_ddot, dtype=np.float64)
        out[f"rtf{i}_complete"] = np.array(bool(ro.done[-1]))
    print(f"  rollout_tensor_full: {len(ros)} rollouts, lengths {[len(r) for r in ros]}, {len(inits)} resets")
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(osp.join(OUT, "diff_extras.npz"), **out)


if __name__ == "__main__":
    if len(sys.argv) > 1:  # only the later additions (leaves the earlier files untouched)
        {"stab": gen_stab_envs, "returns": gen_returns, "pend": lambda: gen_env_trajs("pend", n_cases=6),
         "wrappers": gen_wrappers, "bob-d": lambda: gen_env_trajs("bob-d", n_cases=6), "calm": gen_calm_trajs,
         "fnn64": gen_fnn_closed_loop_envs, "dr": gen_dr_moments, "diff": gen_diff_extras,
         "policies": gen_policy_forward}[sys.argv[1]]()
        sys.exit(0)
    print("trajectories")
    gen_env_trajs("qq-su", n_cases=6)
    gen_env_trajs("qcp-su", n_cases=6)
    gen_env_trajs("qcp-su", n_cases=4, simple_dynamics=True)
    gen_env_trajs("qbb", n_cases=6)
    gen_env_trajs("qbb", n_cases=2, simple_dynamics=True)
    gen_env_trajs("bob", n_cases=6)
    gen_env_trajs("omo", n_cases=6)
    print("closed loop"); gen_fnn_closed_loop()
    print("policies"); gen_policy_forward()
    print("qbb kin"); gen_qbb_kin()
    print("step diff"); gen_step_diff()
    print("dr"); gen_dr_moments()
    print("stab"); gen_stab_envs()
    print("returns"); gen_returns()
    print("pend"); gen_env_trajs("pend", n_cases=6)
    print("bob-d"); gen_env_trajs("bob-d", n_cases=6)
    print("wrappers"); gen_wrappers()
    print("calm"); gen_calm_trajs()
    print("fnn64"); gen_fnn_closed_loop_envs()
    print("diff extras"); gen_diff_extras()
    sz = sum(osp.getsize(osp.join(OUT, f)) for f in os.listdir(OUT))
    print(f"golden size: {sz / 1e6:.2f} MB")
