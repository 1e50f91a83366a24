"""
rollout_tensor_full (pyrado/sampling/rollout.py:640-1016 of the fork) on the CUDA engine: the batched differentiable
rollout that sac_wgrad-style learners sample with.  `bsz` lanes step together through `env.step_diff_state`; a lane
whose episode ends (done, or `env.max_steps` reached) is reset in place (masked reset) and keeps stepping until the step
budget `max_steps` (counted over all lanes) is used up; every step records the Jacobians
    obs_grad[ds, ds + da] = d next_obs / d (obs, act)        rew_grad[1, ds + da] = d rew / d (obs, act)
in `env_infos`, which the reference obtains with five autograd passes per step (rollout.py:832-837).

Here the step is the RK4 kernel (`sre_qcp_diff_forward`), and for the env chains the reference uses this function with --
[ObsActCatWrapper] [ActNormWrapper] QCartPoleSwingUpSim -- all 25 partial derivatives come from ONE launch of the
Jacobian kernel (`sre_qcp_diff_jacobian`); any other chain that offers `step_diff_state` takes the autograd route like
the reference.  Histories stay on the device as [step, lane, ...] tensors and are cut into the reference's
`StepSequence`s (completed episodes in order of completion, then the unfinished ones with `complete=False`) at the end.
"""
from typing import Callable, List, Optional, Union

import numpy as np
import torch
import torch.nn as nn

import simurlacra_b200 as pyrado

from .. import engine as E
from ..environment_wrappers.action_normalization import ActNormWrapper, ObsActCatWrapper
from ..environment_wrappers.base import EnvWrapper
from ..environment_wrappers.utils import all_envs, inner_env
from ..environments.base import Env
from ..environments.sim_base import SimEnv
from ..policies.base import Policy
from .diff import step_jacobians
from .step_sequence import StepSequence


def _kernel_chain(env):
    """(sim, act_scale, act_shift) when the chain is [ObsActCat] [ActNorm] QCartPoleSwingUpSim, else None."""
    from ..environments.pysim.quanser_cartpole import QCartPoleSwingUpSim

    sim = inner_env(env)
    if not isinstance(sim, QCartPoleSwingUpSim):
        return None
    scale, shift = 1.0, 0.0
    for w in [e for e in all_envs(env) if isinstance(e, EnvWrapper)]:
        if isinstance(w, ObsActCatWrapper):
            continue
        if isinstance(w, ActNormWrapper):
            lb, ub = w.wrapped_env.act_space.bounds
            scale, shift = float((ub[0] - lb[0]) / 2), float(lb[0] + (ub[0] - lb[0]) / 2)
            continue
        return None
    return sim, scale, shift


def _autograd_jacobians(obs_next, rew, obs_to, act_to):
    """The reference's route (rollout.py:832-837): one backward pass per output row."""
    rows = [torch.cat(torch.autograd.grad(obs_next[:, i].sum(), [obs_to, act_to], retain_graph=True), dim=1)
            for i in range(obs_next.shape[1])]
    obs_grad = torch.stack(rows, dim=1)
    rew_grad = torch.cat(torch.autograd.grad(rew.sum(), [obs_to, act_to]), dim=1).unsqueeze(1)
    return obs_grad, rew_grad


def rollout_tensor_full(env: Env, policy: Union[nn.Module, Policy, Callable], eval: bool = False,
                        max_steps: Optional[int] = None, reset_kwargs: Optional[dict] = None, render_mode=None,
                        render_step: int = 1, no_reset: bool = False, no_close: bool = False, record_dts: bool = False,
                        stop_on_done: bool = True, seed: Optional[int] = None, sub_seed: Optional[int] = None,
                        sub_sub_seed: Optional[int] = None, bsz: int = 1, num_rollouts: Optional[int] = None,
                        pb=None) -> List[StepSequence]:
    """
    Batched differentiable rollout with masked resets; same signature as the reference (rollout.py:640-658).

    :param max_steps: step budget over ALL lanes (the loop runs `ceil(max_steps / bsz)` iterations)
    :param bsz: number of lanes stepped together
    :param num_rollouts: stop as soon as this many episodes have completed
    :param reset_kwargs: forwarded to the resets; `init_state` may be one state (every reset starts there, like the
                         reference) or an array [k, ds] of states that the resets consume in call order (lanes in order
                         at the start, then the lanes that ended in lane order; engine extension used by the parity test)
    :return: list of `StepSequence`s with `env_infos = {obs_grad, rew_grad}` and `th_ddot`
    """
    if not isinstance(env, Env):
        raise pyrado.TypeErr(given=env, expected_type=Env)
    if not isinstance(eval, bool):
        raise pyrado.TypeErr(given=eval, expected_type=bool)
    if not (isinstance(reset_kwargs, dict) or reset_kwargs is None):
        raise pyrado.TypeErr(given=reset_kwargs, expected_type=dict)
    if not hasattr(env, "step_diff_state"):
        raise pyrado.TypeErr(msg="rollout_tensor_full needs an environment with a differentiable step (step_diff_state)")
    if max_steps is None:
        raise pyrado.ValueErr(msg="rollout_tensor_full needs the step budget max_steps")
    reset_kwargs = dict() if reset_kwargs is None else dict(reset_kwargs)
    if seed is not None:
        pyrado.set_seed(seed, sub_seed, sub_sub_seed)
    sim = inner_env(env)
    dev, dt_ = sim.device, sim.dtype
    ds = sim.state_space.flat_dim
    kc = _kernel_chain(env)
    horizon = env.max_steps

    # ---- initial states of the resets: a fixed state, a list consumed in call order, or the engine's init sampler
    init_arg = reset_kwargs.get("init_state")
    init_list, init_fixed, n_used = None, None, 0
    if init_arg is not None:
        a = torch.as_tensor(np.asarray(init_arg, dtype=np.float64), dtype=dt_, device=dev)
        if a.dim() == 2:
            init_list = a
        else:
            init_fixed = a.reshape(1, ds)
    if reset_kwargs.get("domain_param") is not None:
        env.domain_param = reset_kwargs["domain_param"]
    base_seed = pyrado.seed_hash(seed or 0, sub_seed or 0, sub_sub_seed or 0) if seed is not None else pyrado.device_seed()
    reset_round = [0]

    def fresh_states(n_lanes: int, which: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Initial states for a reset of the lanes in `which` (all lanes if None); [bsz, ds], rows of other lanes unused."""
        nonlocal n_used
        if init_fixed is not None:
            return init_fixed.expand(n_lanes, ds)
        if init_list is not None:
            idx = torch.arange(n_lanes, device=dev) if which is None else torch.nonzero(which).reshape(-1)
            out = torch.zeros(n_lanes, ds, dtype=dt_, device=dev)
            take = init_list[n_used: n_used + idx.numel()]
            if take.shape[0] < idx.numel():
                raise pyrado.ValueErr(msg="reset_kwargs['init_state'] holds fewer states than the rollout resets")
            out[idx] = take
            n_used += idx.numel()
            return out
        cfg = sim._make_cfg()
        reset_round[0] += 1
        return E.sample_init(cfg, None, n_lanes, seed=base_seed, env_offset=0, episode=reset_round[0], device=dev)

    obs = torch.zeros(bsz, ds, dtype=dt_, device=dev) if no_reset else fresh_states(bsz).clone()
    th_ddot = torch.zeros(bsz, dtype=dt_, device=dev)
    steps_ep = torch.zeros(bsz, dtype=torch.int64, device=dev)
    if isinstance(policy, Policy):
        policy.reset(**(dict(domain_param=env.domain_param) if isinstance(env, SimEnv) else dict()))
        policy.eval() if eval else policy.train()
    pol_dev = next(iter(policy.parameters()), torch.zeros(0)).device if isinstance(policy, nn.Module) else None

    H = dict(obs=[], act=[], rew=[], thdd=[], obs_grad=[], rew_grad=[], done=[], nobs=[], nthdd=[])
    curr_steps, n_done_total = 0, 0
    while curr_steps < max_steps:
        if bool(torch.isnan(obs).any()):
            raise pyrado.ValueErr(msg="At least one observation value is NaN!")
        with torch.no_grad():
            o_pol = obs.to(torch.get_default_dtype())
            act = policy(o_pol.to(pol_dev) if pol_dev is not None else o_pol)
        act = act.detach().to(device=dev, dtype=dt_).reshape(bsz, -1)
        if kc is not None:
            with torch.no_grad():
                nobs, rew, done, info = env.step_diff_state(obs, act, th_ddot)
            volts = act * kc[1] + kc[2]
            jac = step_jacobians(kc[0]._diff_cfg(dt_), kc[0]._diff_params(bsz, dt_), obs, volts, th_ddot)  # [B, 5, 5] w.r.t. volts
            jac = jac.clone()
            jac[:, :, ds:] *= kc[1]  # chain rule through the de-normalisation
            obs_grad, rew_grad = jac[:, :ds, :], jac[:, ds:, :]
        else:
            o_req, a_req = obs.clone().requires_grad_(True), act.clone().requires_grad_(True)
            nobs, rew, done, info = env.step_diff_state(o_req, a_req, th_ddot)
            obs_grad, rew_grad = _autograd_jacobians(nobs, rew, o_req, a_req)
            nobs, rew = nobs.detach(), rew.detach()
        steps_ep += 1
        done = done.bool() | (steps_ep >= horizon)
        n_thdd = info["th_ddot"].detach().clone()
        for k, v in (("obs", obs), ("act", act), ("rew", rew), ("thdd", th_ddot), ("obs_grad", obs_grad),
                     ("rew_grad", rew_grad), ("done", done), ("nobs", nobs), ("nthdd", n_thdd)):
            H[k].append(v.detach().clone() if k in ("obs", "thdd") else v.detach())
        # ---- masked reset of the lanes that ended
        if no_reset:
            obs, th_ddot = nobs, n_thdd
        else:
            if init_list is not None:
                new = fresh_states(bsz, done) if bool(done.any()) else obs
            else:
                new = fresh_states(bsz)
            obs = torch.where(done.unsqueeze(1), new, nobs)
            th_ddot = torch.where(done, torch.zeros_like(n_thdd), n_thdd)
            steps_ep = torch.where(done, torch.zeros_like(steps_ep), steps_ep)
        curr_steps += bsz
        if num_rollouts is not None:
            n_done_total += int(done.sum())
            if n_done_total >= num_rollouts:
                break
    if not no_close:
        env.close()
    return _cut_rollouts(env, sim, H, obs, th_ddot, bsz, num_rollouts, pb)


def _cut_rollouts(env, sim, H, last_obs, last_thdd, bsz, num_rollouts, pb) -> List[StepSequence]:
    """[step, lane, ...] histories -> the reference's list: completed episodes in order of completion (step-major, lane-
    minor, rollout.py:860-895), then the unfinished ones lane by lane with `complete=False` (:967-985)."""
    to_np = lambda xs: torch.stack(xs).double().cpu().numpy()
    obs, act, rew, thdd, og, rg, nobs, nthdd = (to_np(H[k]) for k in ("obs", "act", "rew", "thdd", "obs_grad", "rew_grad",
                                                                      "nobs", "nthdd"))
    done = torch.stack(H["done"]).cpu().numpy()
    S = done.shape[0]
    lo_a, up_a = env.act_space.bounds
    info = dict(env_name=env.name, env_spec=env.spec)
    if isinstance(sim, SimEnv):
        info["domain_param"] = env.domain_param
    dt = env.dt
    out, start = [], np.zeros(bsz, dtype=np.int64)

    def make(i, a, b, final_obs, final_thdd, complete):
        sl = slice(a, b)
        observations = np.concatenate([obs[sl, i], final_obs[None]], 0)
        ro = StepSequence(rewards=rew[sl, i], observations=observations, actions=act[sl, i],
                          actions_applied=np.clip(act[sl, i], lo_a, up_a), states=observations.copy(),
                          time=np.arange(b - a + 1) * dt, rollout_info=info,
                          env_infos=dict(obs_grad=og[sl, i], rew_grad=rg[sl, i]),
                          th_ddot=np.concatenate([thdd[sl, i], np.atleast_1d(final_thdd)]), complete=complete)
        if pb is not None:
            pb.update(1 if num_rollouts is not None else len(ro))
        return ro

    for t in range(S):
        for i in np.flatnonzero(done[t]):
            out.append(make(i, start[i], t + 1, nobs[t, i], nthdd[t, i], True))
            start[i] = t + 1
            if num_rollouts is not None and len(out) >= num_rollouts:
                return out
    fo, ft = last_obs.double().cpu().numpy(), last_thdd.double().cpu().numpy()
    for i in range(bsz):
        if start[i] < S:
            out.append(make(i, start[i], S, fo[i], ft[i], False))
    return out
