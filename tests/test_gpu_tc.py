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


@pytest.mark.parametrize("n_obs,n_act", [(6, 1), (5, 1), (8, 2), (4, 1), (2, 1)])
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
