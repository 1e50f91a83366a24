"""
CPU-side checks of the boundary: the C-ABI library loads without a GPU, exports every symbol include/sre.h
declares, the ctypes structs have the C layout, host-only entry points agree with the oracle, argument errors
come back as codes (no exceptions cross the ABI), and the product package never imports the oracle.
"""
import ctypes as C
import os.path as osp
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def L():
    from simurlacra_b200 import _lib

    if not osp.exists(_lib.LIB_PATH):
        import __graft_entry__

        __graft_entry__.build()
    _lib.lib()
    return _lib


def test_every_declared_symbol_is_exported(L):
    hdr = open(osp.join(ROOT, "include", "sre.h")).read()
    declared = set(re.findall(r"^(?:int|const char \*)\s*\*?(sre_\w+)\(", hdr, flags=re.M))
    assert declared, "no declarations parsed"
    assert declared == set(L.SIGNATURES), declared ^ set(L.SIGNATURES)
    out = subprocess.run(["nm", "-D", "--defined-only", L.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (sre_\w+)", out))
    assert declared <= exported, declared - exported
    for name in declared:
        getattr(L.lib(), name)


def test_struct_layout_matches_c(L, tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "sre.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu",'
                   "sizeof(sre_env_cfg),sizeof(sre_task),sizeof(sre_policy),sizeof(sre_rollout_io),sizeof(sre_distr),"
                   "offsetof(sre_env_cfg,base_params),offsetof(sre_rollout_io,traj));}")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", osp.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(L.EnvCfg), C.sizeof(L.Task), C.sizeof(L.Policy), C.sizeof(L.RolloutIO), C.sizeof(L.Distr),
            L.EnvCfg.base_params.offset, L.RolloutIO.traj.offset]
    assert got == want


def test_host_entry_points_agree_with_oracle(L, orc):
    import simurlacra_b200.engine as E

    for kind in range(5):
        d, o = E.dims(kind), orc.dims(kind)
        assert all(d[k] == o[k] for k in ("ds", "do", "da", "dp", "nh"))
        for lp in (False, True):
            np.testing.assert_array_equal(E.nominal_params(kind, lp), orc.nominal_params(kind, lp))
        lay = E.record_layout(kind)
        assert lay["width"] == d["do"] + 2 * d["da"] + 1 + d["ds"] + d["nh"]
        t = E.default_task(kind)
        e = orc.OracleEnv(kind, 0.01, 10)
        # same default reward on a probe state
        s = np.linspace(0.1, 0.4, d["ds"])
        e2 = orc.OracleEnv(kind, 0.01, 10)
        e2.set_task(list(t.state_des)[: d["ds"]], np.array(list(t.Q)[: d["ds"] ** 2]), np.array(list(t.R)[: d["da"] ** 2]),
                    t.rew_kind, t.min_rew, t.radial_mask, t.final_flags, t.final_factor)
        assert e.step_rew(s, np.zeros(d["da"])) == e2.step_rew(s, np.zeros(d["da"]))


def test_argument_errors_are_codes(L):
    import simurlacra_b200.engine as E

    lib = L.lib()
    assert lib.sre_env_dims(99, None, None, None, None, None, None) == -2
    assert b"unknown env kind" in lib.sre_last_error()
    cfg = E.make_cfg(L.QQ, 0.01, 10)
    cfg.dt = -1.0  # Env.__init__ raises ValueErr for dt < 0 (environments/base.py:55-60)
    assert lib.sre_observe(C.byref(cfg), C.c_void_p(8), C.c_void_p(8), 1, None) == -1
    cfg = E.make_cfg(L.QQ, 0.01, 10)
    pol = E.make_policy(L.POL_FNN, None, sizes=[6, 16, 16, 1])
    io = L.RolloutIO()
    io.n_envs, io.n_rows = 4, 11
    assert lib.sre_rollout(C.byref(cfg), C.byref(pol), C.byref(io), None) == -4  # weights NULL
    cfg.kind = L.OMO
    assert lib.sre_qcp_diff_jacobian(C.byref(cfg), None, C.c_void_p(8), None, C.c_void_p(8), 1, C.c_void_p(8), None) == -2
    with pytest.raises(L.EngineError):
        L.check(-1, "probe")


def test_product_never_imports_the_oracle():
    pkg = osp.join(ROOT, "simurlacra_b200")
    out = subprocess.run(["grep", "-rlE", r"(import|from)\s+oracle|sre_oracle|orc_", pkg, "--include=*.py", "--include=*.cu",
                          "--include=*.cuh", "--include=*.h"], capture_output=True, text=True).stdout
    assert out.strip() == "", out
