"""
GPU test of the no-allocation entry point sre_qcp_diff_grad_static (include/sre.h; no counterpart in the reference):
forward + adjoint of the differentiable cart-pole step with the tape in a
file-scope static device array.  Called through the C-ABI (ctypes) with caller-owned inputs and outputs and compared
with the existing pair sre_qcp_diff_forward / sre_qcp_diff_backward on two calls in a row (the second with other sizes
and inputs, so that stale tape contents of the first call would show).  The same kernels run on the same values, so the
results are required to be bit-identical -- tighter than the 1e-8 / 2e-3 bars of tests/test_gpu_diff.py.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _case(E, L, dtype, n, horizon, seed, with_g_states):
    cfg = E.make_cfg(L.QCP, 1 / 50.0, horizon, dtype=dtype, simple_dynamics=False, long_pole=True, wild_init=2)
    g = torch.Generator(device="cpu").manual_seed(seed)
    s0 = (torch.rand(n, 4, generator=g, dtype=torch.float64) - 0.5) * torch.tensor([0.4, 6.0, 1.0, 4.0], dtype=torch.float64)
    act = (torch.rand(horizon, n, generator=g, dtype=torch.float64) - 0.5) * 16.0  # some beyond the +-6 V clamp
    thdd0 = (torch.rand(n, generator=g, dtype=torch.float64) - 0.5)
    g_rew = torch.rand(horizon, n, generator=g, dtype=torch.float64)
    g_states = torch.randn(horizon + 1, n, 4, generator=g, dtype=torch.float64) * 0.1 if with_g_states else None
    to = lambda x: None if x is None else x.to(dtype).to(DEV).contiguous()
    return cfg, to(s0), to(thdd0), to(act), to(g_states), to(g_rew)


def _existing(E, cfg, s0, thdd0, act, g_states, g_rew):
    states, thdd, rew, done = E.qcp_diff_forward(cfg, None, s0, act, thdd0=thdd0)
    g_s0, g_act, g_thdd0 = E.qcp_diff_backward(cfg, None, states, thdd, act, g_states, g_rew)
    return rew, done, g_s0, g_act, g_thdd0


def _static(L, cfg, s0, thdd0, act, g_states, g_rew):
    H, N = act.shape
    dt = act.dtype
    # outputs are the caller's; pre-filled with a poison value so that an element the call skipped would be seen
    rew = torch.full((H, N), float("nan"), dtype=dt, device=DEV)
    done = torch.full((H, N), 255, dtype=torch.uint8, device=DEV)
    g_s0 = torch.full((N, 4), float("nan"), dtype=dt, device=DEV)
    g_act = torch.full((H, N), float("nan"), dtype=dt, device=DEV)
    g_thdd0 = torch.full((N,), float("nan"), dtype=dt, device=DEV)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = L.lib().sre_qcp_diff_grad_static(C.byref(cfg), None, _ptr(s0), _ptr(thdd0), _ptr(act), _ptr(g_states), _ptr(g_rew),
                                          None, N, H, _ptr(rew), _ptr(done), _ptr(g_s0), _ptr(g_act), _ptr(g_thdd0),
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, L.lib().sre_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return rew, done, g_s0, g_act, g_thdd0


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_static_tape_entry_point_equals_forward_backward_pair_twice(dtype):
    import simurlacra_b200.engine as E
    from simurlacra_b200 import _lib as L

    torch.cuda.set_device(0)
    # two calls in a row on the shared tape: a larger one first, then a smaller one with other inputs and no g_states
    for n, horizon, seed, with_gs in ((1000, 120, 1, True), (333, 57, 2, False)):
        cfg, s0, thdd0, act, g_states, g_rew = _case(E, L, dtype, n, horizon, seed, with_gs)
        got = _static(L, cfg, s0, thdd0, act, g_states, g_rew)
        want = _existing(E, cfg, s0, thdd0, act, g_states, g_rew)
        for name, a, b in zip(("rew", "done", "g_s0", "g_act", "g_thdd0"), got, want):
            a, b = a.cpu().numpy(), b.cpu().numpy()
            assert np.isfinite(a.astype(np.float64)).all(), name
            np.testing.assert_array_equal(a, b, err_msg=f"{name} (n={n}, H={horizon}, {dtype})")


def test_static_tape_rejects_sizes_beyond_the_limit():
    import simurlacra_b200.engine as E
    from simurlacra_b200 import _lib as L

    torch.cuda.set_device(0)
    cfg, s0, thdd0, act, g_states, g_rew = _case(E, L, torch.float32, 8, 4, 3, False)
    out = torch.empty(64, dtype=torch.float32, device=DEV)
    limit = 32 << 20
    n, horizon = 1 << 16, 600  # 601 * 65536 * 20 B = 788 MB of tape
    assert (horizon + 1) * n * 20 > limit
    rc = L.lib().sre_qcp_diff_grad_static(C.byref(cfg), None, _ptr(s0), None, _ptr(act), None, None, None, n, horizon, None,
                                          None, _ptr(out), _ptr(out), None, None)
    assert rc == L.E_ARG if hasattr(L, "E_ARG") else rc == -1
    msg = L.lib().sre_last_error().decode()
    assert "SRE_STATIC_TAPE_BYTES" in msg and "65536" in msg, msg
    torch.cuda.synchronize()
