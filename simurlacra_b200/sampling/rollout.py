"""
rollout() (pyrado/sampling/rollout.py:63-343) on the CUDA engine.

Fused path (the hot one): when every wrapper of the chain fuses into the kernels and the policy has a CUDA forward
pass, the WHOLE rollout -- reset, [domain randomization], and the loop {observe (+noise) -> policy -> [ActNorm] ->
SimPyEnv.step} until done -- is ONE launch of the persistent-horizon kernel (`sre_rollout`), for all `num_envs`
environments of the batch at once; the time-major record buffer is then cut into the reference's `StepSequence`s.

Generic path: any other env chain / policy (custom wrappers, arbitrary callables, `no_reset`) runs the reference's loop;
each `env.step` is a launch of the step kernel, the policy is evaluated as a torch module in between.

A batch env (`num_envs > 1`) returns a list with one StepSequence per environment.
"""
import gc
import time
from typing import Callable, List, Optional, Union

import numpy as np
import torch
import torch.nn as nn

import simurlacra_b200 as pyrado

from .. import _lib as L
from .. import engine as E
from ..environment_wrappers.base import EnvWrapper
from ..environment_wrappers.domain_randomization import DomainRandWrapperBuffer, DomainRandWrapperLive
from ..environment_wrappers.utils import all_envs, inner_env
from ..environments.base import Env
from ..environments.pysim.base import Fusion, SimPyEnv
from ..environments.sim_base import SimEnv
from ..policies.base import Policy
from ..utils.data_types import RenderMode
from .step_sequence import StepSequence

HIDDEN_FIELD = {L.QCP: "th_ddot", L.QBB: "plate_angs"}  # name of the hidden-state column(s) in a StepSequence


class CompiledChain:
    """An env chain reduced to what the rollout kernel needs."""

    def __init__(self, sim: SimPyEnv, fusion: Fusion, live_randomizer, buffer_wrappers, obs_layout=None):
        self.sim, self.fusion, self.live_randomizer, self.buffer_wrappers = sim, fusion, live_randomizer, buffer_wrappers
        # with fused ObsPartial / StateAugmentation wrappers: (number of observation entries the record holds,
        # parameter indices, nominal values, fixed) -- how the policy's observation is rebuilt from a record
        self.obs_layout = obs_layout


def _compile_step_wrappers(sim: SimPyEnv, chain: list, fusion: Fusion):
    """The stateful / shape-changing step wrappers of `chain` (outermost first) as the `wrap` dict of engine.make_cfg, for
    the orders the persistent kernel implements (include/sre.h, sre_step_wrappers); False if this chain has to step.
    Observation side, walked from the innermost wrapper outwards with the provenance of every entry."""
    from scipy import signal

    from ..environment_wrappers.action_delay import ActDelayWrapper
    from ..environment_wrappers.action_noise import GaussianActNoiseWrapper
    from ..environment_wrappers.action_normalization import ActNormWrapper
    from ..environment_wrappers.downsampling import DownsamplingWrapper
    from ..environment_wrappers.observation_partial import ObsPartialWrapper
    from ..environment_wrappers.observation_velfilter import ObsVelFiltWrapper
    from ..environment_wrappers.state_augmentation import StateAugmentationWrapper

    kinds = (ActDelayWrapper, DownsamplingWrapper, ObsPartialWrapper, ObsVelFiltWrapper, StateAugmentationWrapper)
    sw = [w for w in chain if isinstance(w, kinds)]
    if not sw:
        return None, None
    if len({type(w) for w in sw}) != len(sw) or fusion.noise_std is not None or fusion.obs_norm is not None:
        return False, None  # two of a kind, or observation noise / normalisation whose place in the order matters
    depth = {type(w): i for i, w in enumerate(chain)}  # larger = closer to the simulation
    d_obs, wrap = sim._dims["do"], {}
    delay, down = depth.get(ActDelayWrapper), depth.get(DownsamplingWrapper)
    norm = depth.get(ActNormWrapper)
    if (delay is not None or down is not None) and any(isinstance(w, GaussianActNoiseWrapper) for w in chain):
        return False, None
    if delay is not None:
        w = chain[delay]
        d = w._delay
        if isinstance(d, np.ndarray) and d.size != 1:
            return False, None  # one delay per env: host path
        d = int(w.delay)
        if d > L.MAX_DELAY:
            return False, None
        inner = norm is not None and delay > norm
        if inner and down is not None and down > delay:
            return False, None  # hold below an inner delay: the kernel holds before the de-normalisation
        wrap.update(act_delay=d, act_delay_inner=inner, delay_first=(down is not None and delay < down))
    if down is not None:
        w = chain[down]
        if w._obs_filter_fcn is not None or not 1 <= int(w.factor) <= L.MAX_DOWN:
            return False, None
        wrap["downsample"] = int(w.factor)
    src, seen_down = [("o", i) for i in range(d_obs)], False
    for w in [w for w in chain[::-1] if isinstance(w, kinds)]:
        if isinstance(w, ObsVelFiltWrapper):
            if seen_down or len(w._b) != 2 or int(w.mask_pos.sum()) > L.MAX_VF or len(w.mask_vel) != len(src):
                return False, None  # filter above the averaging, or not first order
            vel = [src[i] for i in np.flatnonzero(w.mask_vel)]
            if any(k != "o" for k, *_ in vel):
                return False, None
            wrap["velfilt"] = (list(np.flatnonzero(w.mask_pos)), [v[1] for v in vel], w._b[0], w._b[1], w._a[1],
                               float(signal.lfilter_zi(w.b, w.a)[0]))
        elif isinstance(w, DownsamplingWrapper):
            if any(k == "a" for k, *_ in src):
                return False, None  # averaging over appended parameters carries the last episode's values
            seen_down = True
        elif isinstance(w, ObsPartialWrapper):
            if len(w.keep_mask) != len(src):
                return False, None
            src = [src[i] for i in np.flatnonzero(w.keep_mask)]
        elif isinstance(w, StateAugmentationWrapper):
            names = list(sim._param_names)
            if any(k not in names for k in w._params):
                return False, None
            src = src + [("a", names.index(k), float(v), bool(w.fixed)) for k, v in zip(w._params, w._nominal)]
    obs_part = [e[1] for e in src if e[0] == "o"]
    aug = [e for e in src if e[0] == "a"]
    if len(aug) > L.MAX_AUG or not obs_part and not aug:
        return False, None
    if obs_part != list(range(d_obs)):
        wrap["obs_keep"] = obs_part
    if aug:
        wrap["augment"] = ([e[1] for e in aug], [e[2] for e in aug], aug[0][3])
    layout = (len(obs_part), [e[1] for e in aug], [e[2] for e in aug], bool(aug and aug[0][3]))
    return wrap, layout


def compile_chain(env: Env) -> Optional[CompiledChain]:
    """(inner SimPyEnv, Fusion, live randomizer, buffer DR wrappers) if the whole chain fuses, else None."""
    sim = inner_env(env)
    if not isinstance(sim, SimPyEnv):
        return None
    fusion, live, buffers = Fusion(), None, []
    chain = [e for e in all_envs(env) if isinstance(e, EnvWrapper)]
    for i, w in enumerate(chain):
        below = chain[i + 1:]
        if not w._fuse_rollout(fusion, below):
            return None
        if isinstance(w, DomainRandWrapperLive):
            if live is not None:
                return None  # two live randomizers: order-dependent, take the generic path
            live = w.randomizer
        elif isinstance(w, DomainRandWrapperBuffer):
            buffers.append(w)
    if sim._custom_init_space and type(sim.init_space).__name__ != "BoxSpace":
        return None
    wrap, layout = _compile_step_wrappers(sim, chain, fusion)
    if wrap is False:
        return None
    fusion.wrap = wrap
    return CompiledChain(sim, fusion, live, buffers, layout)


def _rollout_info(env: Env, sim, params_row: Optional[np.ndarray]) -> dict:
    info = dict(env_name=env.name, env_spec=env.spec)
    if isinstance(sim, SimEnv):
        if params_row is not None and isinstance(sim, SimPyEnv):
            dp = {k: float(params_row[i]) for i, k in enumerate(sim._param_names)}
            for w in [e for e in all_envs(env) if isinstance(e, EnvWrapper)][::-1]:
                w._set_wrapper_domain_param(dp)
            info["domain_param"] = dp
        else:
            info["domain_param"] = env.domain_param
    return info


class LazyRolloutInfo(dict):
    """`rollout_info` of one rollout of a batch: `env_name` / `env_spec` are shared, the `domain_param` dict of the
    rollout's own (randomized) parameters is built from its row of the kernel's parameter output when it is first asked
    for -- a sampler call returns tens of thousands of rollouts, most consumers read a handful of these dicts."""

    __slots__ = ("_row", "_names", "_wrappers")

    def __init__(self, base: dict, row, names, wrappers):
        super().__init__(base)
        self._row, self._names, self._wrappers = row, names, wrappers

    def _materialise(self):
        if self._row is not None:
            dp = {k: float(self._row[i]) for i, k in enumerate(self._names)}
            for w in self._wrappers:
                w._set_wrapper_domain_param(dp)
            self._row = None
            super().__setitem__("domain_param", dp)

    def __getitem__(self, key):
        if key == "domain_param":
            self._materialise()
        return super().__getitem__(key)

    def get(self, key, default=None):
        if key == "domain_param":
            self._materialise()
        return super().get(key, default)

    def __contains__(self, key):
        return key == "domain_param" and self._row is not None or super().__contains__(key)

    def keys(self):
        self._materialise()
        return super().keys()

    def items(self):
        self._materialise()
        return super().items()

    def values(self):
        self._materialise()
        return super().values()

    def __iter__(self):
        self._materialise()
        return super().__iter__()

    def __len__(self):
        self._materialise()
        return super().__len__()

    def __reduce__(self):
        self._materialise()
        return (dict, (dict(self),))


def sequences_from_records(env: Env, sim: SimPyEnv, res: dict, dt: float, data_format: str = "numpy",
                           env_ids: Optional[np.ndarray] = None, obs_layout=None) -> List[StepSequence]:
    """Cut the first episode of every lane out of the kernel's time-major record buffer (engine extension).
    'torch': every field of every StepSequence is a strided VIEW of the [T, N, W] device buffer (zero copy);
    'numpy': the used part of the buffer crosses to the host in ONE copy and the fields are views of that array.
    Per rollout only the container object is created."""
    length = res["length"].cpu().numpy()
    n = length.shape[0]
    ids = np.arange(n) if env_ids is None else np.asarray(env_ids)
    if len(ids) == 0:
        return []
    t_max = int(length[ids].max()) + 1
    lay, d = E.record_layout(sim.kind), sim._dims
    if data_format == "numpy":
        traj = res["traj"][:t_max].cpu().numpy().astype(np.float64, copy=False)  # one D2H copy
        time_all = np.arange(t_max + 1) * dt
    else:
        traj = res["traj"][:t_max]
        time_all = torch.arange(t_max + 1, dtype=torch.get_default_dtype()) * dt
    o0, a0, p0, r0, s0, h0 = lay["obs"], lay["act"], lay["act_app"], lay["rew"], lay["state"], lay["hidden"]
    do, da, ds, nh = d["do"], d["da"], d["ds"], d["nh"]
    aug_cols = None
    if obs_layout is not None:  # fused ObsPartial / StateAugmentation: the record holds the kept entries; parameters are appended
        do, aug_idx, aug_nom, aug_fixed = obs_layout
        if aug_idx:
            prm = res["params"][:, aug_idx] if not aug_fixed else torch.as_tensor(aug_nom, dtype=res["params"].dtype,
                                                                                 device=res["params"].device).expand(n, -1)
            aug_cols = prm.double().cpu().numpy() if data_format == "numpy" else prm.to(res["traj"].dtype)
    flags_last = res["row_flag"][:t_max].cpu().numpy()
    params = res["params"].double().cpu().numpy()
    names = sim._param_names if isinstance(sim, SimPyEnv) else ()
    wrappers = [e for e in all_envs(env) if isinstance(e, EnvWrapper)][::-1]
    base_info = dict(env_name=env.name, env_spec=env.spec)
    hname = HIDDEN_FIELD.get(sim.kind, "hidden")
    out = []
    # tens of thousands of small containers are created below and none of them is garbage: the cyclic collector would
    # only rescan them again and again (it took 40 % of this loop)
    gc_was_on = gc.isenabled()
    gc.disable()
    for i in ids:
        ln = int(length[i])
        col = traj[:, i]  # [t_max, W] view
        fields = dict(rewards=col[:ln, r0], observations=col[:ln + 1, o0:o0 + do], actions=col[:ln, a0:a0 + da],
                      actions_applied=col[:ln, p0:p0 + da], states=col[:ln + 1, s0:s0 + ds])
        if nh:
            fields[hname] = col[:ln + 1, h0] if nh == 1 else col[:ln + 1, h0:h0 + nh]
        if aug_cols is not None:
            o = fields["observations"]
            fields["observations"] = (np.concatenate((o, np.broadcast_to(aug_cols[i], (ln + 1, aug_cols.shape[1]))), axis=1)
                                      if data_format == "numpy" else torch.cat((o, aug_cols[i].expand(ln + 1, -1)), dim=1))
        ro = StepSequence.from_views(fields, ln, time_all[:ln + 1], LazyRolloutInfo(base_info, params[i], names, wrappers),
                                     True, data_format)
        ro.env_done = bool(flags_last[ln - 1, i] == L.ROW_DONE)  # False if the row budget (not the env) ended the episode
        out.append(ro)
    if gc_was_on:
        gc.enable()
    return out


def _fused_rollout(env, cc: CompiledChain, desc, max_steps, reset_kwargs, stop_on_done, data_format="numpy"):
    sim = cc.sim
    init_state = reset_kwargs.get("init_state")
    domain_param = reset_kwargs.get("domain_param")
    if domain_param is None:
        for w in cc.buffer_wrappers:  # a buffer wrapper hands out its next parameter set(s)
            n = sim.num_envs
            sets = [w._next_set() for _ in range(n)] if isinstance(w.buffer, list) and n > 1 else [w._next_set()]
            domain_param = sets[0] if len(sets) == 1 else {
                k: np.array([float(np.asarray(s[k])) for s in sets]) for k in sets[0] if np.asarray(sets[0][k]).size == 1}
    distr = None
    if domain_param is not None:
        env.domain_param = domain_param  # through the chain: wrappers take their own entries
        cc = compile_chain(env)           # a wrapper parameter (e.g. obs_noise_std) may have changed
    elif cc.live_randomizer is not None:
        distr = cc.live_randomizer.table(sim._param_names)
    cfg = sim._make_cfg(cc.fusion)
    n, d = sim.num_envs, sim._dims
    horizon = int(sim.max_steps) if np.isfinite(sim.max_steps) else None
    if horizon is None:
        raise pyrado.ValueErr(msg="The fused rollout needs a finite max_steps")
    init = None
    if init_state is not None:
        init_state = np.asarray(init_state, dtype=np.float64) if not isinstance(init_state, torch.Tensor) else init_state
        last = init_state.shape[-1]
        if last not in (d["ds"], d["dinit"]):
            raise pyrado.ShapeErr(given=init_state, expected_match=sim.state_space)
        init = sim._to_dev(init_state, last)
        if init.shape[0] == 1 and n > 1:
            init = init.expand(n, -1).contiguous()
    sim._episode += 1
    res = E.rollout(cfg, desc, n, horizon + 1, device=sim.device, params=sim._params, distr=distr, init=init,
                    seed=pyrado.device_seed(), env_offset=sim.env_offset, episode=sim._episode,
                    stop_on_done=stop_on_done, auto_reset=False, record=True)
    status = res["status"]
    if n == 1 or bool((status != 0).any()):
        bad = int(status.max())
        if bad & (L.ST_NAN_OBS | L.ST_NAN_STATE):
            raise pyrado.ValueErr(msg="At least one observation value is NaN!")
        if bad & L.ST_NAN_ACT:
            raise pyrado.ValueErr(msg="At least one action value is NaN!")
    # leave the env where the rollout ended (env.state, curr_step), like the stepping loop would
    sim._alloc()
    ln = res["length"].long()
    idx = torch.arange(n, device=sim.device)
    sim._state.copy_(res["state"][ln, idx])
    if d["nh"] > 0:
        sim._hidden[:, : d["nh"]].copy_(res["hidden"][ln, idx])
    sim._curr_step_t.copy_(res["length"])
    sim._curr_step = int(res["length"][0]) if n == 1 else int(res["length"].max())
    if distr is not None:
        sim.set_param_tensor(res["params"])
        first = res["params"][0].double().cpu().numpy()
        for i, k in enumerate(sim._param_names):
            sim._domain_param[k] = float(first[i])
    return sequences_from_records(env, sim, res, sim.dt, data_format, obs_layout=cc.obs_layout)


def _state_copy(env):
    """rollout.py:237 `env.state.copy()`: a bare batched SimPyEnv hands out its LIVE device tensor, which the step
    kernel overwrites in place -- without the copy every recorded state would alias the final one."""
    s = env.state
    return s.clone() if isinstance(s, torch.Tensor) else np.array(s, copy=True)


def _hidden_of(sim):
    if isinstance(sim, SimPyEnv) and sim._dims["nh"] > 0 and sim.hidden_state is not None:
        return sim.hidden_state.clone()
    return None


def _generic_rollout(env, policy, eval, reset_kwargs, no_reset, stop_on_done, record_dts):
    """The reference's loop (rollout.py:139-312); every env.step is a step-kernel launch."""
    sim = inner_env(env)
    n = getattr(sim, "num_envs", 1)
    batch = n > 1
    obs = (np.zeros(env.obs_space.shape) if not batch else torch.zeros(n, *env.obs_space.shape, device=sim.device)) \
        if no_reset else env.reset(**reset_kwargs)
    info0 = dict(env_name=env.name, env_spec=env.spec)
    if isinstance(sim, SimEnv):
        info0["domain_param"] = env.domain_param
    if isinstance(policy, Policy):
        policy.reset(**(dict(domain_param=env.domain_param) if isinstance(env, SimEnv) else dict()))
        policy.eval() if eval else policy.train()
    hist = dict(obs=[], act=[], act_app=[], rew=[], state=[], hidden=[], done=[], dt_policy=[], dt_step=[])
    finished = torch.zeros(n, dtype=torch.bool, device=sim.device) if batch else False
    pol_dev = next(iter(policy.parameters()), torch.zeros(0)).device if isinstance(policy, nn.Module) else None
    steps = 0
    while steps < env.max_steps and not (bool(finished.all()) if batch else (finished and stop_on_done)):
        t0 = time.time()
        obs_t = obs if isinstance(obs, torch.Tensor) else torch.from_numpy(np.asarray(obs))
        if bool(torch.isnan(obs_t).any()):
            raise pyrado.ValueErr(msg="At least one observation value is NaN!")
        obs_to = obs_t.to(dtype=torch.get_default_dtype())
        if pol_dev is not None:
            obs_to = obs_to.to(pol_dev)
        with torch.no_grad():
            act_to = policy(obs_to)
        if bool(torch.isnan(act_to).any()):
            raise pyrado.ValueErr(msg="At least one action value is NaN!")
        act = act_to.detach().to(sim.device) if batch else act_to.detach().cpu().numpy()
        t1 = time.time()
        state = _state_copy(env)
        hid = _hidden_of(sim)
        obs_next, rew, done, _ = env.step(act)
        act_app = env.limit_act(act)
        t2 = time.time()
        for k, v in (("obs", obs), ("act", act), ("act_app", act_app), ("rew", rew), ("state", state), ("hidden", hid),
                     ("done", done), ("dt_policy", t1 - t0), ("dt_step", t2 - t1)):
            hist[k].append(v)
        obs = obs_next
        steps += 1
        finished = (finished | done) if batch and stop_on_done else (done if not batch else finished)
    hist["obs"].append(obs)
    hist["state"].append(_state_copy(env))
    hist["hidden"].append(_hidden_of(sim))
    if steps == 0:
        raise pyrado.ShapeErr(msg="StepSequence must not be empty!")
    hname = HIDDEN_FIELD.get(getattr(sim, "kind", None), "hidden")

    def seq_for(i, ln):
        pick = (lambda v: v) if not batch else (lambda v: v[i].double().cpu().numpy() if isinstance(v, torch.Tensor) else v)
        data = dict(actions_applied=[pick(a) for a in hist["act_app"][:ln]], states=[pick(s) for s in hist["state"][:ln + 1]])
        if hist["hidden"][0] is not None:
            hs = [h[i if batch else 0].double().cpu().numpy() for h in hist["hidden"][:ln + 1]]
            data[hname] = [h[0] if h.shape[0] == 1 else h for h in hs]
        if record_dts:
            data["dts_policy"], data["dts_step"] = hist["dt_policy"][:ln], hist["dt_step"][:ln]
        return StepSequence(rewards=[float(pick(r)) for r in hist["rew"][:ln]],
                            observations=[pick(o) for o in hist["obs"][:ln + 1]],
                            actions=[pick(a) for a in hist["act"][:ln]], time=np.arange(ln + 1) * env.dt,
                            rollout_info=info0, complete=True, **data)

    if not batch:
        return [seq_for(0, steps)]
    done_mat = torch.stack(hist["done"]).cpu().numpy()  # [steps, n]
    out = []
    for i in range(n):
        hit = np.flatnonzero(done_mat[:, i])
        out.append(seq_for(i, int(hit[0]) + 1 if len(hit) and stop_on_done else steps))
    return out


def rollout(env: Env, policy: Union[nn.Module, Policy, Callable], eval: bool = False, max_steps: Optional[int] = None,
            reset_kwargs: Optional[dict] = None, render_mode: RenderMode = RenderMode(), render_step: int = 1,
            no_reset: bool = False, no_close: bool = False, record_dts: bool = False, stop_on_done: bool = True,
            seed: Optional[int] = None, sub_seed: Optional[int] = None, sub_sub_seed: Optional[int] = None,
            data_format: str = "numpy") -> Union[StepSequence, List[StepSequence]]:
    """
    Perform a rollout (i.e. sample a trajectory) in the given environment using the given policy.  Same signature as
    the reference (rollout.py:63-78); rendering arguments are accepted and ignored (no graphics stack on this box).

    :return: one `StepSequence` for a single env, a list with one `StepSequence` per env for a batch env
    """
    if not isinstance(env, Env):
        raise pyrado.TypeErr(given=env, expected_type=Env)
    if not isinstance(eval, bool):
        raise pyrado.TypeErr(given=eval, expected_type=bool)
    if not (isinstance(reset_kwargs, dict) or reset_kwargs is None):
        raise pyrado.TypeErr(given=reset_kwargs, expected_type=dict)
    if not isinstance(render_mode, RenderMode):
        raise pyrado.TypeErr(given=render_mode, expected_type=RenderMode)
    if max_steps is not None:
        env.max_steps = max_steps
    if seed is not None:
        pyrado.set_seed(seed, sub_seed, sub_sub_seed)  # host generators, like the reference
        # Device streams: rollout `sub_sub_seed` of sampling round `sub_seed` is lane `sub_sub_seed` of the Philox
        # stream keyed seed_hash(seed, sub_seed) -- the key ParallelRolloutSampler gives to its lane of that index,
        # so that sequential rollouts equal the sampler's (Pyrado/tests/test_sampling.py:570-587)
        sim0 = inner_env(env)
        if isinstance(sim0, SimPyEnv):
            pyrado.set_device_seed(pyrado.seed_hash(seed, sub_seed, 0))
            sim0.env_offset = (sub_sub_seed or 0) * sim0.num_envs
            sim0._episode = -1  # the reset below advances it to episode 0
    reset_kwargs = dict() if reset_kwargs is None else reset_kwargs

    cc = None if (no_reset or record_dts) else compile_chain(env)
    desc = None
    if cc is not None and isinstance(policy, Policy) and np.isfinite(env.max_steps):
        desc = policy.engine_desc(cc.sim.device)
    if cc is not None and desc is not None:
        policy.reset(**(dict(domain_param=env.domain_param) if isinstance(env, SimEnv) else dict()))
        policy.eval() if eval else policy.train()
        ros = _fused_rollout(env, cc, desc, max_steps, reset_kwargs, stop_on_done, data_format)
    else:
        ros = _generic_rollout(env, policy, eval, reset_kwargs, no_reset, stop_on_done, record_dts)
    if not no_close:
        env.close()
    return ros[0] if getattr(inner_env(env), "num_envs", 1) == 1 else ros


def after_rollout_query(env: Env, policy: Policy, rollout: StepSequence) -> tuple:
    """Non-interactive stand-in for the reference's console prompt (rollout.py:1209-1299): always "continue"."""
    return False, None, None
