"""
Thin functional layer between torch tensors and the C-ABI (``include/sre.h``).  Everything above it (the
Pyrado-style classes in ``environments/``, ``sampling/`` ...) goes through these calls; nothing here computes on the
host.  torch is used for device memory and streams only.
"""
import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib as L

KIND_BY_NAME = {"qq-su": L.QQ, "qcp-su": L.QCP, "qbb": L.QBB, "bob": L.BOB, "omo": L.OMO, "pend": L.PEND}


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def dims(kind: int) -> dict:
    v = [C.c_int() for _ in range(6)]
    L.check(L.lib().sre_env_dims(kind, *[C.byref(x) for x in v]), "sre_env_dims")
    return dict(zip(("ds", "do", "da", "dp", "nh", "dinit"), (x.value for x in v)))


def nominal_params(kind: int, long_pole: bool = False) -> np.ndarray:
    out = np.zeros(L.MAX_DP)
    L.check(L.lib().sre_nominal_params(kind, int(long_pole), out.ctypes.data_as(C.c_void_p)), "sre_nominal_params")
    return out[: dims(kind)["dp"]].copy()


def record_layout(kind: int) -> dict:
    v = [C.c_int() for _ in range(6)]
    L.check(L.lib().sre_record_layout(kind, *[C.byref(x) for x in v]), "sre_record_layout")
    d = dict(zip(("obs", "act", "act_app", "rew", "state", "hidden"), (x.value for x in v)))
    d["width"] = L.lib().sre_record_width(kind)
    return d


def default_task(kind: int) -> L.Task:
    t = L.Task()
    L.check(L.lib().sre_default_task(kind, C.byref(t)), "sre_default_task")
    return t


def make_task(kind: int, state_des=None, Q=None, R=None, rew_kind=None, min_rew=None, radial_idcs=None,
              radial_mod=None, final_flags=None, final_factor=None) -> L.Task:
    """Default task of the env with overrides (the reference's ``task_args``)."""
    d = dims(kind)
    t = default_task(kind)
    if state_des is not None:
        for i, v in enumerate(np.asarray(state_des, dtype=np.float64).reshape(-1)):
            t.state_des[i] = v
    if Q is not None:
        Q = np.asarray(Q, dtype=np.float64).reshape(d["ds"], d["ds"])
        for i, v in enumerate(Q.reshape(-1)):
            t.Q[i] = v
    if R is not None:
        R = np.asarray(R, dtype=np.float64).reshape(d["da"], d["da"])
        for i, v in enumerate(R.reshape(-1)):
            t.R[i] = v
    if rew_kind is not None:
        t.rew_kind = rew_kind
    if min_rew is not None:
        t.min_rew = min_rew
    if radial_idcs is not None:
        t.radial_mask = sum(1 << int(i) for i in radial_idcs)
        if radial_mod is not None:
            mod = np.ones(len(radial_idcs)) * np.asarray(radial_mod, dtype=np.float64)
            for i, m in zip(radial_idcs, mod):
                t.radial_mod[int(i)] = m
    if final_flags is not None:
        t.final_flags = final_flags
    if final_factor is not None:
        t.final_factor = final_factor
    return t


def make_cfg(kind: int, dt: float, max_steps: int = 0, dtype=torch.float32, simple_dynamics: bool = False,
             long_pole: bool = False, wild_init: int = 0, task: Optional[L.Task] = None, base_params=None,
             act_norm: bool = False, noise_std=None, noise_mean=None, init_box=None, flags: int = 0,
             obs_norm=None, env_act_noise=None, wrap: Optional[dict] = None) -> L.EnvCfg:
    """`obs_norm` = (lb, ub) of ObsNormWrapper; `env_act_noise` = (position, std, mean) of GaussianActNoiseWrapper with
    position L.ENV_ACT_NOISE_OUTER (outside ActNormWrapper) or L.ENV_ACT_NOISE_INNER; `wrap` = the stateful step-path
    wrappers of the persistent rollout (sre_step_wrappers): keys act_delay, act_delay_inner, delay_first, downsample,
    obs_keep (indices kept by ObsPartialWrapper), velfilt = (pos state indices, vel obs indices, b0, b1, a1, zi),
    augment = (parameter indices, nominal values, fixed)."""
    cfg = L.EnvCfg()
    cfg.kind = kind
    cfg.dtype = L.F64 if dtype == torch.float64 else L.F32
    cfg.flags = (L.F_SIMPLE_DYN if simple_dynamics else 0) | int(flags)
    cfg.wild_init = wild_init
    cfg.dt = float(dt)
    cfg.max_steps = int(max_steps) if max_steps and np.isfinite(max_steps) else 0
    cfg.task = task if task is not None else default_task(kind)
    bp = nominal_params(kind, long_pole) if base_params is None else np.asarray(base_params, dtype=np.float64)
    for i, v in enumerate(bp):
        cfg.base_params[i] = v
    cfg.act_norm = int(act_norm)
    if noise_std is not None:
        cfg.obs_noise = 1
        for i, v in enumerate(np.asarray(noise_std, dtype=np.float64).reshape(-1)):
            cfg.noise_std[i] = v
        if noise_mean is not None:
            for i, v in enumerate(np.asarray(noise_mean, dtype=np.float64).reshape(-1)):
                cfg.noise_mean[i] = v
    if obs_norm is not None:
        cfg.obs_norm = 1
        for i, v in enumerate(np.asarray(obs_norm[0], dtype=np.float64).reshape(-1)):
            cfg.obs_lb[i] = v
        for i, v in enumerate(np.asarray(obs_norm[1], dtype=np.float64).reshape(-1)):
            cfg.obs_ub[i] = v
    if env_act_noise is not None:
        cfg.env_act_noise = int(env_act_noise[0])
        for i, v in enumerate(np.asarray(env_act_noise[1], dtype=np.float64).reshape(-1)):
            cfg.env_act_noise_std[i] = v
        for i, v in enumerate(np.asarray(env_act_noise[2], dtype=np.float64).reshape(-1)):
            cfg.env_act_noise_mean[i] = v
    if wrap:
        w = cfg.wrap
        w.act_delay = int(wrap.get("act_delay", 0))
        w.act_delay_inner = int(bool(wrap.get("act_delay_inner", False)))
        w.delay_first = int(bool(wrap.get("delay_first", False)))
        w.downsample = int(wrap.get("downsample", 0))
        if wrap.get("obs_keep") is not None:
            w.obs_keep_mask = sum(1 << int(i) for i in wrap["obs_keep"])
        if wrap.get("velfilt") is not None:
            pos, vel, b0, b1, a1, zi = wrap["velfilt"]
            w.vf_n = len(pos)
            for i, (p_, v_) in enumerate(zip(pos, vel)):
                w.vf_pos[i], w.vf_vel[i] = int(p_), int(v_)
            w.vf_b0, w.vf_b1, w.vf_a1, w.vf_zi = float(b0), float(b1), float(a1), float(zi)
        if wrap.get("augment") is not None:
            idx, nominal, fixed = wrap["augment"]
            w.n_aug, w.aug_fixed = len(idx), int(bool(fixed))
            for i, (p_, v_) in enumerate(zip(idx, nominal)):
                w.aug_param[i], w.aug_nominal[i] = int(p_), float(v_)
    if init_box is not None:
        cfg.has_init_box = 1
        for i, v in enumerate(np.asarray(init_box[0], dtype=np.float64).reshape(-1)):
            cfg.init_box_lo[i] = v
        for i, v in enumerate(np.asarray(init_box[1], dtype=np.float64).reshape(-1)):
            cfg.init_box_up[i] = v
    return cfg


def cfg_dtype(cfg: L.EnvCfg):
    return torch.float64 if cfg.dtype == L.F64 else torch.float32


def make_distr_table(entries: Sequence[dict]):
    """entries: dicts(param_index, kind in {'normal','uniform','bernoulli'}, a, b[, c, clip_lo, clip_up, roundint])"""
    arr = (L.Distr * max(len(entries), 1))()
    kinds = dict(normal=L.DISTR_NORMAL, uniform=L.DISTR_UNIFORM, bernoulli=L.DISTR_BERNOULLI)
    for i, e in enumerate(entries):
        arr[i].param_index = int(e["param_index"])
        arr[i].kind = kinds[e["kind"]] if isinstance(e["kind"], str) else int(e["kind"])
        arr[i].a, arr[i].b, arr[i].c = float(e["a"]), float(e["b"]), float(e.get("c", 0.0))
        arr[i].clip_lo = float(e.get("clip_lo", -np.inf))
        arr[i].clip_up = float(e.get("clip_up", np.inf))
        arr[i].roundint = int(e.get("roundint", False))
    arr._n = len(entries)
    return arr


def make_policy(kind: int, weights: Optional[torch.Tensor] = None, sizes=(), feats=(), hidden_nonlin: int = 0,
                output_nonlin: int = 0, cartpole_feats: bool = False, precision: int = L.PREC_TF32X3, act_noise: int = 0,
                act_noise_scale=(), act_noise_mean=(), group_size: int = 0, weights_host=None) -> L.Policy:
    """`group_size` > 0: `weights` is a [n_sets, n_weights] matrix of parameter sets and lane n of a rollout launch
    evaluates set n // group_size (parameter-exploration sampling).
    `precision`: arithmetic of a [in, 64, 64, out] tanh network's hidden layer in the float32 build -- L.PREC_TF32X3
    (default, tcgen05 3 x TF32 split, fp32-level accuracy), L.PREC_FP32 (CUDA-core FMA), L.PREC_TF32 (single pass);
    other shapes and the float64 build ignore it."""
    p = L.Policy()
    set_act_noise(p, act_noise, act_noise_scale, act_noise_mean)
    p.kind = kind
    p.n_layers = max(len(sizes) - 1, 0)
    for i, s in enumerate(sizes):
        p.sizes[i] = int(s)
    p.n_feats = len(feats)
    blocks = []  # a feature is a name / code, or (name, [parameters]) for the parametrised ones (include/sre.h SRE_FEAT_SIG ..)
    for i, f in enumerate(feats):
        if isinstance(f, (tuple, list)):
            p.feat_arg[i] = sum(len(b) for b in blocks)
            blocks.append([float(v) for v in f[1]])
            f = f[0]
        p.feats[i] = L.FEATS[f] if isinstance(f, str) else int(f)
    if blocks:
        flat = [v for b in blocks for v in b]
        if len(flat) > L.MAX_FEAT_PARAMS:
            raise ValueError(f"{len(flat)} feature parameters, the kernel holds {L.MAX_FEAT_PARAMS}")
        p._feat_keep = (C.c_float * len(flat))(*flat)  # HOST array, copied into the launch by the library
        p.feat_params, p.n_feat_params = C.cast(p._feat_keep, C.c_void_p), len(flat)
    p.hidden_nonlin, p.output_nonlin, p.cartpole_feats = hidden_nonlin, output_nonlin, int(cartpole_feats)
    p.precision = int(precision)
    if weights_host is not None and not group_size:  # optional host copy (float32, contiguous, CPU) of the same parameters
        wh = weights_host if isinstance(weights_host, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(weights_host, dtype=np.float32))
        if wh.is_cuda or wh.dtype != torch.float32 or not wh.is_contiguous() or (weights is not None and wh.numel() != weights.numel()):
            raise ValueError("weights_host must be a contiguous float32 CPU tensor with as many entries as weights")
        p._keep_host = wh
        p.weights_host = wh.data_ptr()
    if weights is not None:
        if not (weights.is_cuda and weights.dtype == torch.float32 and weights.is_contiguous()):
            raise ValueError("policy weights must be a contiguous float32 CUDA tensor")
        p._keep = weights
        p.weights = weights.data_ptr()
        p.n_weights = weights.numel()
        if group_size:
            if weights.dim() != 2:
                raise ValueError("grouped policies take a [n_sets, n_weights] weight matrix")
            p.group_size, p.n_groups, p.n_weights = int(group_size), weights.shape[0], weights.shape[1]
    return p


def set_act_noise(p: L.Policy, kind: int, scale=(), mean=()) -> L.Policy:
    """Step-based exploration fused into the rollout kernel: kind L.ACT_NOISE_NORMAL (scale = std) or
    L.ACT_NOISE_UNIFORM (scale = halfspan); `mean` is the noise's own mean (train_mean), zeros if empty."""
    p.act_noise = int(kind)
    for i in range(L.MAX_DA):
        p.act_noise_scale[i] = float(scale[i]) if i < len(scale) else 0.0
        p.act_noise_mean[i] = float(mean[i]) if i < len(mean) else 0.0
    return p


def _chk(t: Optional[torch.Tensor], dtype, name: str):
    if t is None:
        return None
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (no CPU fallback)")
    if t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous {dtype}, got {t.dtype} contiguous={t.is_contiguous()}")
    return t


def stats_init(device) -> torch.Tensor:
    return torch.tensor([0, 0, 0, 0, 0, 0, 1e300, -1e300], dtype=torch.float64, device=device)


def derive(cfg, params: Optional[torch.Tensor], n: int, device=None) -> dict:
    """domain_param setter side effects: spaces and c_max per env."""
    dt_, d = cfg_dtype(cfg), dims(cfg.kind)
    device = params.device if params is not None else device
    out = dict(st_lo=torch.empty(n, d["ds"], dtype=dt_, device=device), st_up=torch.empty(n, d["ds"], dtype=dt_, device=device),
               act_lo=torch.empty(n, d["da"], dtype=dt_, device=device), act_up=torch.empty(n, d["da"], dtype=dt_, device=device),
               c_max=torch.empty(n, dtype=dt_, device=device))
    L.check(L.lib().sre_derive(C.byref(cfg), _ptr(_chk(params, dt_, "params")), n, _ptr(out["st_lo"]), _ptr(out["st_up"]),
                               _ptr(out["act_lo"]), _ptr(out["act_up"]), _ptr(out["c_max"]), _stream()), "sre_derive")
    return out


def sample_params(cfg, distr, params: torch.Tensor, seed: int, env_offset: int = 0, episode: int = 0):
    _chk(params, cfg_dtype(cfg), "params")
    L.check(L.lib().sre_sample_params(C.byref(cfg), C.cast(distr, C.c_void_p), distr._n, seed, env_offset, episode,
                                      _ptr(params), params.shape[0], _stream()), "sre_sample_params")
    return params


def sample_init(cfg, params: Optional[torch.Tensor], n: int, seed: int, env_offset: int = 0, episode: int = 0,
                device=None) -> torch.Tensor:
    dt_, d = cfg_dtype(cfg), dims(cfg.kind)
    device = params.device if params is not None else device
    init = torch.empty(n, d["dinit"], dtype=dt_, device=device)
    L.check(L.lib().sre_sample_init(C.byref(cfg), _ptr(_chk(params, dt_, "params")), seed, env_offset, episode, _ptr(init),
                                    n, _stream()), "sre_sample_init")
    return init


def reset(cfg, params, init: torch.Tensor, state, hidden, curr_step, obs=None, mask=None, seed=0, env_offset=0,
          episode=0):
    dt_ = cfg_dtype(cfg)
    _chk(init, dt_, "init"); _chk(state, dt_, "state"); _chk(hidden, dt_, "hidden"); _chk(curr_step, torch.int32, "curr_step")
    L.check(L.lib().sre_reset(C.byref(cfg), _ptr(_chk(params, dt_, "params")), _ptr(init), init.shape[1],
                              _ptr(_chk(mask, torch.uint8, "mask")), _ptr(state), _ptr(hidden), _ptr(curr_step),
                              _ptr(_chk(obs, dt_, "obs")), seed, env_offset, episode, state.shape[0], _stream()), "sre_reset")


def step(cfg, params, act, state, hidden, curr_step, obs, rew, done, status=None, seed=0, env_offset=0, episode=0):
    dt_ = cfg_dtype(cfg)
    for t, nm in ((act, "act"), (state, "state"), (hidden, "hidden"), (obs, "obs"), (rew, "rew")):
        _chk(t, dt_, nm)
    L.check(L.lib().sre_step(C.byref(cfg), _ptr(_chk(params, dt_, "params")), _ptr(act), _ptr(state), _ptr(hidden),
                             _ptr(_chk(curr_step, torch.int32, "curr_step")), _ptr(obs), _ptr(rew),
                             _ptr(_chk(done, torch.uint8, "done")), _ptr(_chk(status, torch.int32, "status")), seed,
                             env_offset, episode, state.shape[0], _stream()), "sre_step")


def observe(cfg, state: torch.Tensor) -> torch.Tensor:
    dt_, d = cfg_dtype(cfg), dims(cfg.kind)
    obs = torch.empty(state.shape[0], d["do"], dtype=dt_, device=state.device)
    L.check(L.lib().sre_observe(C.byref(cfg), _ptr(_chk(state, dt_, "state")), _ptr(obs), state.shape[0], _stream()),
            "sre_observe")
    return obs


def step_rew(cfg, params, state: torch.Tensor, act: torch.Tensor) -> torch.Tensor:
    dt_ = cfg_dtype(cfg)
    rew = torch.empty(state.shape[0], dtype=dt_, device=state.device)
    L.check(L.lib().sre_step_rew(C.byref(cfg), _ptr(_chk(params, dt_, "params")), _ptr(_chk(state, dt_, "state")),
                                 _ptr(_chk(act, dt_, "act")), _ptr(rew), state.shape[0], _stream()), "sre_step_rew")
    return rew


def task_step_rew(task: L.Task, ds: int, da: int, c_max: float, state: torch.Tensor, act: torch.Tensor) -> torch.Tensor:
    """Reward of a free-standing task for a batch: state [N, ds], act [N, da] (same dtype, float32 or float64)."""
    dt_ = state.dtype
    rew = torch.empty(state.shape[0], dtype=dt_, device=state.device)
    L.check(L.lib().sre_task_step_rew(C.byref(task), L.F64 if dt_ == torch.float64 else L.F32, ds, da, float(c_max),
                                      _ptr(_chk(state, dt_, "state")), _ptr(_chk(act, dt_, "act")), _ptr(rew), state.shape[0],
                                      _stream()), "sre_task_step_rew")
    return rew


def policy_forward(pol: L.Policy, obs: torch.Tensor, n_act: int) -> torch.Tensor:
    _chk(obs, torch.float32, "obs")
    act = torch.empty(obs.shape[0], n_act, dtype=torch.float32, device=obs.device)
    L.check(L.lib().sre_policy_forward(C.byref(pol), _ptr(obs), obs.shape[1], _ptr(act), n_act, obs.shape[0], _stream()),
            "sre_policy_forward")
    return act


def rollout(cfg, pol: L.Policy, n_envs: int, n_rows: int, *, device, params=None, distr=None, init=None, actions=None,
            seed: int = 0, env_offset: int = 0, episode: int = 0, stop_on_done: bool = True, auto_reset: bool = False,
            record: bool = True, want_stats: bool = True, traj: Optional[torch.Tensor] = None,
            row_flag: Optional[torch.Tensor] = None, traj_env_offset: Optional[int] = None) -> dict:
    """One launch of the persistent-horizon kernel. Returns device tensors (views into `traj` for the fields).
    `traj_env_offset`: `traj` [T, M, W] / `row_flag` [T, M] are WIDER than this launch (M >= offset + n_envs, e.g. the
    learner's buffer for all ranks, possibly peer-mapped); lane n writes column offset + n and the returned field views
    are the launch's own columns."""
    dt_, d = cfg_dtype(cfg), dims(cfg.kind)
    lay = record_layout(cfg.kind)
    io = L.RolloutIO()
    io.n_envs, io.n_rows = n_envs, n_rows
    io.seed, io.env_offset, io.episode = seed, env_offset, episode
    io.stop_on_done, io.auto_reset = int(stop_on_done), int(auto_reset)
    if distr is not None:
        io.distr, io.n_distr = C.cast(distr, C.c_void_p), distr._n
    io.params = _ptr(_chk(params, dt_, "params"))
    if init is not None:
        io.init, io.n_init_cols = _ptr(_chk(init, dt_, "init")), init.shape[1]
    io.actions = _ptr(_chk(actions, dt_, "actions"))
    res = {}
    if record:
        if traj is None:
            traj = torch.empty(n_rows, n_envs, lay["width"], dtype=dt_, device=device)
        if row_flag is None:
            row_flag = torch.empty(n_rows, n_envs, dtype=torch.uint8, device=device)
        io.traj, io.row_flag = _ptr(_chk(traj, dt_, "traj")), _ptr(_chk(row_flag, torch.uint8, "row_flag"))
        if traj_env_offset is not None:
            if traj.shape[1] != row_flag.shape[1] or traj_env_offset < 0 or traj_env_offset + n_envs > traj.shape[1]:
                raise ValueError("traj / row_flag do not hold columns [offset, offset + n_envs)")
            io.traj_env_stride, io.traj_env_offset = traj.shape[1], int(traj_env_offset)
            traj, row_flag = traj[:, traj_env_offset: traj_env_offset + n_envs], row_flag[:, traj_env_offset: traj_env_offset + n_envs]
        res["traj"], res["row_flag"] = traj, row_flag
        res["obs"] = traj[..., lay["obs"]: lay["obs"] + d["do"]]
        res["act"] = traj[..., lay["act"]: lay["act"] + d["da"]]
        res["act_app"] = traj[..., lay["act_app"]: lay["act_app"] + d["da"]]
        res["rew"] = traj[..., lay["rew"]]
        res["state"] = traj[..., lay["state"]: lay["state"] + d["ds"]]
        res["hidden"] = traj[..., lay["hidden"]: lay["hidden"] + d["nh"]]
    res["ret"] = torch.zeros(n_envs, dtype=dt_, device=device)
    res["length"] = torch.zeros(n_envs, dtype=torch.int32, device=device)
    res["params"] = torch.empty(n_envs, d["dp"], dtype=dt_, device=device)
    res["status"] = torch.zeros(n_envs, dtype=torch.int32, device=device)
    io.ret, io.length, io.params_out, io.status = _ptr(res["ret"]), _ptr(res["length"]), _ptr(res["params"]), _ptr(res["status"])
    if want_stats:
        res["stats"] = stats_init(device)
        io.stats = _ptr(res["stats"])
    L.check(L.lib().sre_rollout(C.byref(cfg), C.byref(pol), C.byref(io), _stream()), "sre_rollout")
    return res


def returns(rew: torch.Tensor, row_flag: torch.Tensor, gamma: float, values: Optional[torch.Tensor] = None,
            lamb: float = 0.95, out_ret: Optional[torch.Tensor] = None, out_adv: Optional[torch.Tensor] = None):
    """Discounted returns-to-go (and, with `values`, generalised advantage estimates) of every episode of every lane of
    a time-major buffer, on the device (sre_returns).  `rew` [T, N] may be a strided view (e.g. the `rew` column of the
    record buffer); `row_flag` uint8 [T, N]; `values` [T, N] value estimates of every row's observation.
    Returns (ret [T, N], adv [T, N] or None); rows without a step hold 0.  `out_ret` / `out_adv` reuse caller buffers."""
    if rew.dim() != 2 or row_flag.shape != rew.shape or not rew.is_cuda:
        raise ValueError("rew and row_flag must be CUDA tensors of the same [T, N] shape")
    if rew.dtype not in (torch.float32, torch.float64):
        raise ValueError("rew must be float32 or float64")
    T, N = rew.shape
    flag = _chk(row_flag.contiguous(), torch.uint8, "row_flag")
    out_ret = torch.empty(T, N, dtype=rew.dtype, device=rew.device) if out_ret is None else _chk(out_ret, rew.dtype, "out_ret")
    if out_ret.shape != rew.shape:
        raise ValueError("out_ret must have the shape of rew")
    if values is None:
        out_adv = None
    if values is not None:
        if values.shape != rew.shape:
            raise ValueError("values must have the shape of rew")
        values = _chk(values.to(rew.dtype).contiguous(), rew.dtype, "values")
        out_adv = torch.empty(T, N, dtype=rew.dtype, device=rew.device) if out_adv is None else _chk(out_adv, rew.dtype, "out_adv")
        if out_adv.shape != rew.shape:
            raise ValueError("out_adv must have the shape of rew")
    L.check(L.lib().sre_returns(L.F64 if rew.dtype == torch.float64 else L.F32, rew.data_ptr(), rew.stride(0), rew.stride(1),
                                _ptr(flag), _ptr(values), float(gamma), float(lamb), _ptr(out_ret), _ptr(out_adv), T, N,
                                _stream()), "sre_returns")
    return out_ret, out_adv


def qcp_diff_forward(cfg, params, s0, act, thdd0=None):
    """act [H, N] inner-env units before the clamp. Returns states [H+1,N,4], thdd [H+1,N], rew [H,N], done [H,N]."""
    dt_ = cfg_dtype(cfg)
    H, N = act.shape
    dev = s0.device
    states = torch.empty(H + 1, N, 4, dtype=dt_, device=dev)
    thdd = torch.empty(H + 1, N, dtype=dt_, device=dev)
    rew = torch.empty(H, N, dtype=dt_, device=dev)
    done = torch.empty(H, N, dtype=torch.uint8, device=dev)
    L.check(L.lib().sre_qcp_diff_forward(C.byref(cfg), _ptr(_chk(params, dt_, "params")), _ptr(_chk(s0, dt_, "s0")),
                                         _ptr(_chk(thdd0, dt_, "thdd0")), _ptr(_chk(act, dt_, "act")), N, H, _ptr(states),
                                         _ptr(thdd), _ptr(rew), _ptr(done), _stream()), "sre_qcp_diff_forward")
    return states, thdd, rew, done


def qcp_diff_backward(cfg, params, states, thdd, act, g_states=None, g_rew=None, g_thdd=None):
    dt_ = cfg_dtype(cfg)
    H, N = act.shape
    dev = act.device
    g_s0 = torch.empty(N, 4, dtype=dt_, device=dev)
    g_act = torch.empty(H, N, dtype=dt_, device=dev)
    g_thdd0 = torch.empty(N, dtype=dt_, device=dev)
    L.check(L.lib().sre_qcp_diff_backward(C.byref(cfg), _ptr(_chk(params, dt_, "params")), _ptr(_chk(states, dt_, "states")),
                                          _ptr(_chk(thdd, dt_, "thdd")), _ptr(_chk(act, dt_, "act")),
                                          _ptr(_chk(g_states, dt_, "g_states")), _ptr(_chk(g_rew, dt_, "g_rew")),
                                          _ptr(_chk(g_thdd, dt_, "g_thdd")), N, H, _ptr(g_s0), _ptr(g_act), _ptr(g_thdd0),
                                          _stream()), "sre_qcp_diff_backward")
    return g_s0, g_act, g_thdd0


def qcp_diff_jacobian(cfg, params, state, act, thdd=None):
    dt_ = cfg_dtype(cfg)
    N = state.shape[0]
    jac = torch.empty(N, 5, 5, dtype=dt_, device=state.device)
    L.check(L.lib().sre_qcp_diff_jacobian(C.byref(cfg), _ptr(_chk(params, dt_, "params")), _ptr(_chk(state, dt_, "state")),
                                          _ptr(_chk(thdd, dt_, "thdd")), _ptr(_chk(act, dt_, "act")), N, _ptr(jac), _stream()),
            "sre_qcp_diff_jacobian")
    return jac
