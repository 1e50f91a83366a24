"""
The float32 `m_sincos` of the production build (simurlacra_b200/csrc/sre_device.cuh) against float64.

The reference evaluates the dynamics and observations with numpy float64 (`np.sin`, `np.cos`, e.g.
Pyrado/pyrado/environments/pysim/quanser_qube.py:123-149); the fp32 build has a 1e-4 parity budget over 600 steps, so
its trigonometry has to stay at the level of libm's float functions (<= 2 ulp).  The CPU test restates the device
algorithm in numpy float32 (FMAs emulated through float64) with the constants the header holds; the GPU test calls the
observe kernel through the C-ABI.
"""
import os.path as osp
import re

import numpy as np
import pytest

ROOT = osp.dirname(osp.dirname(osp.abspath(__file__)))
f32 = np.float32

# constants of m_sincos(float): 2/pi, the three Cody-Waite terms of pi/2, the sine and cosine polynomials
TWO_OVER_PI = "0.63661977236758138f"
CW = ["-0x1.921fb6p+0f", "0x1.777a5cp-25f", "0x1.ee59dap-50f"]
SIN_C = ["-1.9515295891e-4f", "8.3321608736e-3f", "-1.6666654611e-1f"]
COS_C = ["2.443315711809948e-5f", "-1.388731625493765e-3f", "4.166664568298827e-2f"]


def _val(tok):
    tok = tok.rstrip("f")
    return float.fromhex(tok) if "0x" in tok else float(tok)


def _fma(a, b, c):
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def lean_sincos(x):
    k = lambda tok: np.full_like(x, _val(tok))
    q = np.rint((x * f32(_val(TWO_OVER_PI))).astype(f32)).astype(f32)
    i = q.astype(np.int64)
    r = _fma(q, k(CW[0]), x)
    r = _fma(q, k(CW[1]), r)
    r = _fma(q, k(CW[2]), r)
    z = (r * r).astype(f32)
    ps = _fma(_fma(k(SIN_C[0]), z, k(SIN_C[1])), z, k(SIN_C[2]))
    ps = _fma((ps * z).astype(f32), r, r)
    pc = _fma(_fma(k(COS_C[0]), z, k(COS_C[1])), z, k(COS_C[2]))
    pc = _fma((pc * z).astype(f32), z, _fma(k("-0.5"), z, k("1.0")))
    odd = (i & 1) == 1
    s, c = np.where(odd, pc, ps), np.where(odd, ps, pc)
    return np.where((i & 2) != 0, -s, s), np.where(((i + 1) & 2) != 0, -c, c)


def test_header_holds_the_constants_this_test_restates():
    src = open(osp.join(ROOT, "simurlacra_b200", "csrc", "sre_device.cuh")).read()
    body = src[src.index("SRE_DEV void m_sincos(float x"):]
    body = body[:body.index("#else")]
    for tok in [TWO_OVER_PI] + CW + SIN_C + COS_C:
        assert tok in body, tok
    assert not re.search(r"\bsincosf\(", body)  # no libm call, no slow path


@pytest.mark.parametrize("span", [3.2, 100.0, 5000.0, 1e5])
def test_lean_sincos_within_two_ulp_of_float64(span):
    rng = np.random.default_rng(int(span))
    x = rng.uniform(-span, span, 1_000_000).astype(f32)
    x[:9] = np.array([0.0, np.pi / 2, -np.pi / 2, np.pi, -np.pi, 1e-30, -1e-30, 2 * np.pi, 1.5 * np.pi], dtype=f32)
    s, c = lean_sincos(x)
    xs = x.astype(np.float64)
    for got, ref in ((s, np.sin(xs)), (c, np.cos(xs))):
        err = np.abs(got.astype(np.float64) - ref)
        assert err.max() < 1.3e-7  # absolute: about one float32 ulp of 1.0
        ulp = np.spacing(np.maximum(np.abs(ref), 1e-30).astype(f32)).astype(np.float64)
        big = np.abs(ref) > 1e-3  # near the zeros the reduced argument's rounding dominates, the absolute bound holds
        assert (err[big] / ulp[big]).max() < 2.0
    assert np.all(s[:1] == 0.0) and np.all(c[:1] == 1.0)


@pytest.mark.gpu
def test_observe_kernel_trigonometry_matches_float64():
    """QQube observation (sin th, cos th, sin al, cos al, th_dot, al_dot), quanser_qube.py:148-149, fp32 build."""
    import torch

    import simurlacra_b200.engine as E
    from simurlacra_b200 import _lib as L

    cfg = E.make_cfg(L.QQ, 1 / 250.0, 600, dtype=torch.float32)
    rng = np.random.default_rng(0)
    n = 1 << 18
    st = np.zeros((n, 4), dtype=f32)
    st[:, 0] = rng.uniform(-1e4, 1e4, n)
    st[:, 1] = rng.uniform(-7.0, 7.0, n)
    st[:4, 0] = [0.0, np.pi / 2, -np.pi, 1e-30]
    st[:, 2:] = rng.standard_normal((n, 2))
    obs = E.observe(cfg, torch.from_numpy(st).cuda()).cpu().numpy().astype(np.float64)
    ref = np.stack([np.sin(st[:, 0].astype(np.float64)), np.cos(st[:, 0].astype(np.float64)),
                    np.sin(st[:, 1].astype(np.float64)), np.cos(st[:, 1].astype(np.float64))], axis=1)
    assert np.abs(obs[:, :4] - ref).max() < 1.3e-7
    assert np.array_equal(obs[:, 4:], st[:, 2:].astype(np.float64))
    # non-finite angles give NaN like np.sin / sincosf
    bad = torch.tensor([[float("inf"), float("nan"), 0.0, 0.0]], dtype=torch.float32, device="cuda")
    assert torch.isnan(E.observe(cfg, bad)[0, :4]).all()
    # the verification build keeps libm: agreement with numpy float64 to 2 ulp
    cfg64 = E.make_cfg(L.QQ, 1 / 250.0, 600, dtype=torch.float64)
    obs64 = E.observe(cfg64, torch.from_numpy(st.astype(np.float64)).cuda()).cpu().numpy()
    assert np.abs(obs64[:, :4] - ref).max() < 5e-16
