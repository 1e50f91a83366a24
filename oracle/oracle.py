"""
TEST INFRASTRUCTURE ONLY -- ctypes binding of the CPU oracle (`oracle/sre_oracle.c`).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import
this module; the product package `simurlacra_b200` never does.
"""
import ctypes as C
import os
import os.path as osp
import subprocess

import numpy as np

_HERE = osp.dirname(osp.abspath(__file__))
_LIB_PATH = osp.join(_HERE, "_build", "libsre_oracle.so")

QQ, QCP, QBB, BOB, OMO, PEND = 0, 1, 2, 3, 4, 5
KIND_BY_NAME = {"qq-su": QQ, "qcp-su": QCP, "qbb": QBB, "bob": BOB, "omo": OMO, "pend": PEND}
F_SIMPLE_DYN = 1
F_QCP_STAB = 2
F_DISC_ACT = 4
REW_QUADR, REW_EXP, REW_SCALED_EXP = 0, 1, 2
FR_ENABLED, FR_STATE_DEP, FR_TIME_DEP, FR_ALWAYS_POS, FR_ALWAYS_NEG = 1, 2, 4, 8, 16
POL_ACTIONS, POL_FNN, POL_LINEAR, POL_IDLE = 0, 1, 2, 3
FEATS = dict(const=0, identity=1, sign=2, abs=3, squared=4, cubic=5, sin=6, cos=7, sinsin=8, sincos=9)

MAX_DS, MAX_DO, MAX_DA, MAX_DP, MAX_NH = 8, 8, 2, 20, 2


def build(force: bool = False) -> str:
    """Compile the C oracle with the committed Makefile (gcc, a second or two)."""
    src = osp.join(_HERE, "sre_oracle.c")
    if force or not osp.exists(_LIB_PATH) or osp.getmtime(_LIB_PATH) < osp.getmtime(src):
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _LIB_PATH


class Distr(C.Structure):
    _fields_ = [
        ("param_index", C.c_int32),
        ("kind", C.c_int32),
        ("a", C.c_float),
        ("b", C.c_float),
        ("c", C.c_float),
        ("clip_lo", C.c_float),
        ("clip_up", C.c_float),
        ("roundint", C.c_int32),
    ]


class Policy(C.Structure):
    _fields_ = [
        ("kind", C.c_int),
        ("n_layers", C.c_int),
        ("sizes", C.c_int * 9),
        ("n_feats", C.c_int),
        ("feats", C.c_int * 12),
        ("hidden_nonlin", C.c_int),
        ("output_nonlin", C.c_int),
        ("cartpole_feats", C.c_int),
        ("weights", C.POINTER(C.c_float)),
        ("act_noise", C.c_int),
        ("act_noise_scale", C.c_float * 2),
        ("act_noise_mean", C.c_float * 2),
        ("group_size", C.c_long), ("n_weights", C.c_long),
    ]


class RolloutCfg(C.Structure):
    _fields_ = [
        ("kind", C.c_int), ("flags", C.c_int), ("long_pole", C.c_int), ("wild_init", C.c_int),
        ("dt", C.c_double),
        ("max_steps", C.c_long),
        ("has_task", C.c_int),
        ("state_des", C.c_double * MAX_DS), ("Q", C.c_double * (MAX_DS * MAX_DS)), ("R", C.c_double * (MAX_DA * MAX_DA)),
        ("rew_kind", C.c_int), ("min_rew", C.c_double), ("radial_mask", C.c_uint), ("final_flags", C.c_int),
        ("final_factor", C.c_double),
        ("act_norm", C.c_int),
        ("obs_noise", C.c_int),
        ("noise_std", C.c_double * MAX_DO), ("noise_mean", C.c_double * MAX_DO),
        ("params", C.c_void_p),
        ("distr", C.c_void_p), ("n_distr", C.c_int),
        ("init", C.c_void_p),
        ("n_init_cols", C.c_int),
        ("init_box_lo", C.c_void_p), ("init_box_up", C.c_void_p),
        ("actions", C.c_void_p),
        ("seed", C.c_uint64), ("env_offset", C.c_uint64), ("episode", C.c_uint32),
        ("stop_on_done", C.c_int),
        ("obs_norm", C.c_int), ("env_act_noise", C.c_int),
        ("obs_lb", C.c_double * MAX_DO), ("obs_ub", C.c_double * MAX_DO),
        ("an_std", C.c_double * MAX_DA), ("an_mean", C.c_double * MAX_DA),
    ]


class RolloutOut(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("obs", "act", "act_app", "rew", "state", "hidden", "done", "ret", "length", "params_out")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_double, C.c_long, C.c_int, C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_set_params.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_get_params.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_set_task.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_uint,
                                   C.c_int, C.c_double]
        L.orc_get_spaces.argtypes = [C.c_void_p] * 5
        L.orc_set_radial_mod.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_get_cmax.restype = C.c_double
        L.orc_get_cmax.argtypes = [C.c_void_p]
        L.orc_get_state.argtypes = [C.c_void_p] * 3
        L.orc_set_state.argtypes = [C.c_void_p] * 3
        L.orc_curr_step.restype = C.c_long
        L.orc_curr_step.argtypes = [C.c_void_p]
        L.orc_observe.argtypes = [C.c_void_p] * 3
        L.orc_qbb_kin.restype = C.c_double
        L.orc_qbb_kin.argtypes = [C.c_double] * 3
        L.orc_reset.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_step_rew.restype = C.c_double
        L.orc_step_rew.argtypes = [C.c_void_p] * 3
        L.orc_limit_act.argtypes = [C.c_void_p] * 3
        L.orc_step.restype = C.c_int
        L.orc_step.argtypes = [C.c_void_p] * 4
        L.orc_qcp_step_diff_state.restype = C.c_int
        L.orc_qcp_step_diff_state.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_void_p,
                                              C.c_void_p, C.c_void_p]
        L.orc_philox4x32_10.argtypes = [C.c_void_p] * 3
        L.orc_rng_block.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_sample_entry.restype = C.c_float
        L.orc_sample_entry.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32]
        L.orc_sample_params.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_uint32]
        L.orc_sample_init.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32,
                                      C.c_void_p, C.c_void_p]
        L.orc_policy_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orc_rollout_batch.restype = C.c_long
        L.orc_rollout_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_void_p, C.c_int]
        L.orc_nominal_params.argtypes = [C.c_int, C.c_int, C.c_void_p]
        for f in ("orc_ds", "orc_do", "orc_da", "orc_dp", "orc_nh"):
            getattr(L, f).argtypes = [C.c_int]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def dims(kind):
    L = lib()
    return dict(ds=L.orc_ds(kind), do=L.orc_do(kind), da=L.orc_da(kind), dp=L.orc_dp(kind), nh=L.orc_nh(kind))


def nominal_params(kind, long_pole=False):
    out = np.zeros(MAX_DP)
    lib().orc_nominal_params(kind, int(long_pole), _p(out))
    return out[: dims(kind)["dp"]].copy()


def philox(ctr, key):
    ctr = np.asarray(ctr, dtype=np.uint32)
    key = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    lib().orc_philox4x32_10(_p(ctr), _p(key), _p(out))
    return out


def qbb_kin(arm_radius, plate_length, th):
    return lib().orc_qbb_kin(float(arm_radius), float(plate_length), float(th))


def make_distr_table(entries):
    """entries: list of dicts(param_index, kind ('normal'|'uniform'|'bernoulli'), a, b, c, clip_lo, clip_up, roundint)"""
    arr = (Distr * len(entries))()
    kinds = dict(normal=0, uniform=1, bernoulli=2)
    for i, e in enumerate(entries):
        arr[i].param_index = e["param_index"]
        arr[i].kind = kinds[e["kind"]]
        arr[i].a = e["a"]
        arr[i].b = e["b"]
        arr[i].c = e.get("c", 0.0)
        arr[i].clip_lo = e.get("clip_lo", -np.inf)
        arr[i].clip_up = e.get("clip_up", np.inf)
        arr[i].roundint = int(e.get("roundint", False))
    return arr


def make_policy(kind, weights=None, sizes=(), feats=(), hidden_nonlin=0, output_nonlin=0, cartpole_feats=False,
                act_noise=0, act_noise_scale=(), act_noise_mean=(), group_size=0):
    p = Policy()
    p.group_size = int(group_size)
    if group_size and weights is not None:
        p.n_weights = np.asarray(weights).reshape(-1).size // (np.asarray(weights).shape[0] if np.asarray(weights).ndim == 2 else 1)
    p.act_noise = int(act_noise)
    for i, v in enumerate(act_noise_scale):
        p.act_noise_scale[i] = v
    for i, v in enumerate(act_noise_mean):
        p.act_noise_mean[i] = v
    p.kind = kind
    p.n_layers = max(len(sizes) - 1, 0)
    for i, s in enumerate(sizes):
        p.sizes[i] = s
    p.n_feats = len(feats)
    for i, f in enumerate(feats):
        p.feats[i] = FEATS[f] if isinstance(f, str) else f
    p.hidden_nonlin = hidden_nonlin
    p.output_nonlin = output_nonlin
    p.cartpole_feats = int(cartpole_feats)
    if weights is not None:
        w = np.ascontiguousarray(weights, dtype=np.float32)
        p._keep = w
        p.weights = w.ctypes.data_as(C.POINTER(C.c_float))
    return p


def policy_forward(pol, obs, n_act):
    obs = np.ascontiguousarray(obs, dtype=np.float32)
    act = np.zeros(n_act, dtype=np.float32)
    lib().orc_policy_forward(C.byref(pol), _p(obs), obs.size, _p(act), n_act)
    return act


class OracleEnv:
    """One environment of the C oracle with the reference's reset/step semantics (fp64)."""

    def __init__(self, kind, dt, max_steps=0, simple_dynamics=False, long_pole=False, stab=False, disc_act=False):
        self.kind = kind
        self.d = dims(kind)
        self._L = lib()
        self._h = self._L.orc_create(kind, float(dt), int(max_steps),
                                     (F_SIMPLE_DYN if simple_dynamics else 0) | (F_QCP_STAB if stab else 0)
                                     | (F_DISC_ACT if disc_act else 0), int(long_pole))

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.orc_destroy(self._h)
            self._h = None

    def set_params(self, dp):
        dp = np.ascontiguousarray(dp, dtype=np.float64)
        assert dp.size == self.d["dp"]
        self._L.orc_set_params(self._h, _p(dp))

    def get_params(self):
        out = np.zeros(self.d["dp"])
        self._L.orc_get_params(self._h, _p(out))
        return out

    def set_task(self, state_des, Q, R, rew_kind, min_rew=1e-4, radial_mask=0, final_flags=0, final_factor=0.0):
        sd = np.ascontiguousarray(state_des, dtype=np.float64)
        Q = np.ascontiguousarray(Q, dtype=np.float64)
        R = np.ascontiguousarray(R, dtype=np.float64)
        self._L.orc_set_task(self._h, _p(sd), _p(Q), _p(R), rew_kind, min_rew, radial_mask, final_flags, final_factor)

    def set_radial_mod(self, mod):
        m = np.ascontiguousarray(mod, dtype=np.float64)
        assert m.size == self.d["ds"]
        self._L.orc_set_radial_mod(self._h, _p(m))

    def spaces(self):
        ds, da = self.d["ds"], self.d["da"]
        a, b, c, d = np.zeros(ds), np.zeros(ds), np.zeros(da), np.zeros(da)
        self._L.orc_get_spaces(self._h, _p(a), _p(b), _p(c), _p(d))
        return a, b, c, d

    @property
    def c_max(self):
        return self._L.orc_get_cmax(self._h)

    @property
    def state(self):
        s, h = np.zeros(self.d["ds"]), np.zeros(MAX_NH)
        self._L.orc_get_state(self._h, _p(s), _p(h))
        return s

    @property
    def hidden(self):
        s, h = np.zeros(self.d["ds"]), np.zeros(MAX_NH)
        self._L.orc_get_state(self._h, _p(s), _p(h))
        return h[: self.d["nh"]]

    def set_state(self, state, hidden=None):
        s = np.ascontiguousarray(state, dtype=np.float64)
        h = None
        if hidden is not None:
            h = np.zeros(MAX_NH)
            h[: len(np.atleast_1d(hidden))] = hidden
        self._L.orc_set_state(self._h, _p(s), _p(h))

    def reset(self, init_state):
        init = np.ascontiguousarray(init_state, dtype=np.float64)
        obs = np.zeros(self.d["do"])
        self._L.orc_reset(self._h, _p(init), init.size, _p(obs))
        return obs

    def step(self, act):
        act = np.ascontiguousarray(np.atleast_1d(act), dtype=np.float64)
        obs = np.zeros(self.d["do"])
        rew = C.c_double()
        done = self._L.orc_step(self._h, _p(act), _p(obs), C.byref(rew))
        return obs, rew.value, done

    def step_rew(self, state, act):
        s = np.ascontiguousarray(state, dtype=np.float64)
        a = np.ascontiguousarray(np.atleast_1d(act), dtype=np.float64)
        return self._L.orc_step_rew(self._h, _p(s), _p(a))

    def step_diff_state(self, state, act, act_diff, th_ddot):
        s = np.ascontiguousarray(state, dtype=np.float64)
        ns = np.zeros(4)
        rew, thdd = C.c_double(), C.c_double()
        done = self._L.orc_qcp_step_diff_state(self._h, _p(s), float(act), float(act_diff), float(th_ddot), _p(ns),
                                               C.byref(rew), C.byref(thdd))
        return ns, rew.value, bool(done), thdd.value

    def sample_params(self, table, seed, env, episode):
        self._L.orc_sample_params(self._h, C.byref(table), len(table), seed, env, episode)

    def sample_init(self, seed, env, episode, wild_init=0, box=None):
        init = np.zeros(MAX_DS)
        n = C.c_int()
        lo = np.ascontiguousarray(box[0], dtype=np.float64) if box is not None else None
        up = np.ascontiguousarray(box[1], dtype=np.float64) if box is not None else None
        self._L.orc_sample_init(self._h, wild_init, _p(lo), _p(up), seed, env, episode, _p(init), C.byref(n))
        return init[: n.value].copy()


def rollout_batch(kind, dt, max_steps, N, H, *, policy, simple_dynamics=False, long_pole=False, wild_init=0,
                  task=None, act_norm=False, noise_std=None, noise_mean=None, params=None, distr=None, init=None,
                  init_box=None, actions=None, seed=0, env_offset=0, episode=0, stop_on_done=True, record=True,
                  n_threads=0, stab=False, obs_norm=None, env_act_noise=None, disc_act=False):
    """Run N independent rollouts on the CPU oracle. Returns a dict of arrays ([T, N, d] layout) + 'steps'."""
    d = dims(kind)
    cfg = RolloutCfg()
    cfg.kind, cfg.long_pole, cfg.wild_init = kind, int(long_pole), wild_init
    cfg.flags = (F_SIMPLE_DYN if simple_dynamics else 0) | (F_QCP_STAB if stab else 0) | (F_DISC_ACT if disc_act else 0)
    if obs_norm is not None:
        cfg.obs_norm = 1
        for i, v in enumerate(np.asarray(obs_norm[0], dtype=np.float64).reshape(-1)): cfg.obs_lb[i] = v
        for i, v in enumerate(np.asarray(obs_norm[1], dtype=np.float64).reshape(-1)): cfg.obs_ub[i] = v
    if env_act_noise is not None:
        cfg.env_act_noise = int(env_act_noise[0])
        for i, v in enumerate(np.asarray(env_act_noise[1], dtype=np.float64).reshape(-1)): cfg.an_std[i] = v
        for i, v in enumerate(np.asarray(env_act_noise[2], dtype=np.float64).reshape(-1)): cfg.an_mean[i] = v
    cfg.dt, cfg.max_steps = dt, max_steps
    keep = []
    if task is not None:
        cfg.has_task = 1
        sd = np.zeros(MAX_DS); sd[: d["ds"]] = task["state_des"]
        Q = np.ascontiguousarray(task["Q"], dtype=np.float64).reshape(-1)
        R = np.ascontiguousarray(task["R"], dtype=np.float64).reshape(-1)
        for i, v in enumerate(sd): cfg.state_des[i] = v
        for i, v in enumerate(Q): cfg.Q[i] = v
        for i, v in enumerate(R): cfg.R[i] = v
        cfg.rew_kind = task["rew_kind"]
        cfg.min_rew = task.get("min_rew", 1e-4)
        cfg.radial_mask = task.get("radial_mask", 0)
        cfg.final_flags = task.get("final_flags", 0)
        cfg.final_factor = task.get("final_factor", 0.0)
    cfg.act_norm = int(act_norm)
    if noise_std is not None:
        cfg.obs_noise = 1
        for i, v in enumerate(noise_std): cfg.noise_std[i] = v
        if noise_mean is not None:
            for i, v in enumerate(noise_mean): cfg.noise_mean[i] = v

    def arr(a, dtype=np.float64):
        if a is None:
            return None
        a = np.ascontiguousarray(a, dtype=dtype)
        keep.append(a)
        return a.ctypes.data

    cfg.params = arr(params)
    if distr is not None:
        keep.append(distr)
        cfg.distr = C.addressof(distr)
        cfg.n_distr = len(distr)
    if init is not None:
        init = np.ascontiguousarray(init, dtype=np.float64)
        cfg.n_init_cols = init.shape[1]
        cfg.init = arr(init)
    if init_box is not None:
        cfg.init_box_lo, cfg.init_box_up = arr(init_box[0]), arr(init_box[1])
    cfg.actions = arr(actions)
    cfg.seed, cfg.env_offset, cfg.episode, cfg.stop_on_done = seed, env_offset, episode, int(stop_on_done)

    out = RolloutOut()
    res = {}
    if record:
        res["obs"] = np.full((H + 1, N, d["do"]), np.nan)
        res["act"] = np.full((H, N, d["da"]), np.nan)
        res["act_app"] = np.full((H, N, d["da"]), np.nan)
        res["rew"] = np.zeros((H, N))
        res["state"] = np.full((H + 1, N, d["ds"]), np.nan)
        res["hidden"] = np.full((H + 1, N, max(d["nh"], 1)), np.nan)
        res["done"] = np.zeros((H, N), dtype=np.uint8)
        out.obs, out.act, out.act_app = res["obs"].ctypes.data, res["act"].ctypes.data, res["act_app"].ctypes.data
        out.rew, out.state, out.done = res["rew"].ctypes.data, res["state"].ctypes.data, res["done"].ctypes.data
        if d["nh"]:
            out.hidden = res["hidden"].ctypes.data
    res["ret"] = np.zeros(N)
    res["length"] = np.zeros(N, dtype=np.int32)
    res["params"] = np.zeros((N, d["dp"]))
    out.ret, out.length, out.params_out = res["ret"].ctypes.data, res["length"].ctypes.data, res["params"].ctypes.data
    res["steps"] = lib().orc_rollout_batch(C.byref(cfg), C.byref(policy), N, H, C.byref(out), n_threads)
    return res


def returns(rew, flag, gamma, values=None, lamb=0.95):
    """orc_returns on [T, N] arrays: (ret, adv or None)."""
    rew = np.ascontiguousarray(rew, dtype=np.float64)
    flag = np.ascontiguousarray(flag, dtype=np.uint8)
    T, N = rew.shape
    ret = np.zeros((T, N))
    adv = np.zeros((T, N)) if values is not None else None
    val = np.ascontiguousarray(values, dtype=np.float64) if values is not None else None
    f = lib().orc_returns
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_long, C.c_long]
    f(rew.ctypes.data, flag.ctypes.data, val.ctypes.data if val is not None else None, gamma, lamb, ret.ctypes.data,
      adv.ctypes.data if adv is not None else None, T, N)
    return ret, adv

