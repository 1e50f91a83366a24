"""
torch.autograd bindings of the differentiable cart-pole path (quanser_cartpole.py:257-358 `step_diff_state`, the
path sac_wgrad and rollout_tensor_full drive).  Forward = the RK4 step kernel (`sre_qcp_diff_forward`), backward = the
hand-written adjoint kernel (`sre_qcp_diff_backward`); nothing is traced or recomputed by torch.

  QcpDiffStep     one step:   (state [B,4], act_unclamped [B], th_ddot [B]) -> (next_state, rew, done, th_ddot')
  QcpDiffRollout  H steps:    (s0 [B,4], acts [H,B], th_ddot0 [B])          -> (states [H+1,B,4], rew [H,B], done, thdd)
  step_jacobians  per-step Jacobians d(next_state, rew)/d(state, act) that rollout_tensor_full collects
                  (sampling/rollout.py:832-837), in one kernel launch instead of five autograd passes
"""
import torch

from .. import engine as E


class QcpDiffRollout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cfg, params, s0, acts, thdd0):
        dt_ = E.cfg_dtype(cfg)
        s0c, ac = s0.detach().to(dt_).contiguous(), acts.detach().to(dt_).contiguous()
        t0 = None if thdd0 is None else thdd0.detach().to(dt_).contiguous()
        states, thdd, rew, done = E.qcp_diff_forward(cfg, params, s0c, ac, t0)
        ctx.cfg, ctx.params = cfg, params
        ctx.save_for_backward(states, thdd, ac)
        ctx.in_dtypes = (s0.dtype, acts.dtype, None if thdd0 is None else thdd0.dtype)
        ctx.mark_non_differentiable(done)
        return states, rew, done, thdd

    @staticmethod
    def backward(ctx, g_states, g_rew, _g_done, g_thdd):
        states, thdd, ac = ctx.saved_tensors
        dt_ = states.dtype
        prep = lambda g: None if g is None else g.to(dt_).contiguous()
        g_s0, g_act, g_thdd0 = E.qcp_diff_backward(ctx.cfg, ctx.params, states, thdd, ac, prep(g_states), prep(g_rew),
                                                   prep(g_thdd))
        ds, da, dth = ctx.in_dtypes
        return None, None, g_s0.to(ds), g_act.to(da), (None if dth is None else g_thdd0.to(dth))


class QcpDiffStep(torch.autograd.Function):
    """One differentiable step; the horizon-1 case of the rollout kernels."""

    @staticmethod
    def forward(ctx, cfg, params, state, act_unclamped, thdd):
        dt_ = E.cfg_dtype(cfg)
        s, a = state.detach().to(dt_).contiguous(), act_unclamped.detach().to(dt_).reshape(1, -1).contiguous()
        t = thdd.detach().to(dt_).contiguous()
        states, thdd_tape, rew, done = E.qcp_diff_forward(cfg, params, s, a, t)
        ctx.cfg, ctx.params = cfg, params
        ctx.save_for_backward(states, thdd_tape, a)
        ctx.in_dtypes = (state.dtype, act_unclamped.dtype, thdd.dtype)
        done_b = done[0].bool()
        ctx.mark_non_differentiable(done_b)
        return states[1].to(state.dtype), rew[0].to(state.dtype), done_b, thdd_tape[1].to(state.dtype)

    @staticmethod
    def backward(ctx, g_ns, g_rew, _g_done, g_thdd):
        states, thdd_tape, a = ctx.saved_tensors
        dt_ = states.dtype
        n = states.shape[1]
        g_states = torch.zeros(2, n, 4, dtype=dt_, device=states.device)
        if g_ns is not None:
            g_states[1] = g_ns.to(dt_)
        g_r = torch.zeros(1, n, dtype=dt_, device=states.device) if g_rew is None else g_rew.to(dt_).reshape(1, n).contiguous()
        g_t = None
        if g_thdd is not None:
            g_t = torch.zeros(2, n, dtype=dt_, device=states.device)
            g_t[1] = g_thdd.to(dt_)
        g_s0, g_act, g_thdd0 = E.qcp_diff_backward(ctx.cfg, ctx.params, states, thdd_tape, a, g_states, g_r, g_t)
        ds, da, dth = ctx.in_dtypes
        return None, None, g_s0.to(ds), g_act[0].to(da), g_thdd0.to(dth)


def step_jacobians(cfg, params, state: torch.Tensor, act_unclamped: torch.Tensor, thdd=None) -> torch.Tensor:
    """[B, 5, 5]: rows (s'_0..3, rew), columns (s_0..3, act) -- `obs_grad` and `rew_grad` of rollout_tensor_full."""
    dt_ = E.cfg_dtype(cfg)
    return E.qcp_diff_jacobian(cfg, params, state.detach().to(dt_).contiguous(),
                               act_unclamped.detach().to(dt_).reshape(-1).contiguous(),
                               None if thdd is None else thdd.detach().to(dt_).contiguous())
